"""Drop-in for ``diff_gaussian_rasterization`` (rmurai0610/diff-gaussian-rasterization-w-pose@43e21bf,
reference requirements.txt:22) -- "variant B" of the operator, used by loop-closure registration
through ``src/gsr/renderer.py:15-18,59-73,131-142``.

    color, radii, depth, opacity, n_touched = GaussianRasterizer(raster_settings=s)(
        means3D=..., means2D=..., opacities=..., shs=... | colors_precomp=...,
        scales=..., rotations=... | cov3D_precomp=..., theta=cam_rot_delta, rho=cam_trans_delta)

``opacity`` (1,H,W) = 1 - T_final; ``n_touched`` (P,) int32 = #pixels where the Gaussian was blended
while the transmittance stayed above 0.5.  ``theta.grad`` / ``rho.grad`` (3,) receive the camera-pose
gradient of the left perturbation T_CW <- Exp([rho, theta]) T_CW (reference
src/utils/pose_utils.py:76-93), reduced on the device to one 6-vector instead of upstream's (P,6)
buffer + Python sum.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from eags_raster import _lib
from eags_raster.rasterize import rasterize_gaussians as _rasterize
from gaussian_rasterizer import mark_visible


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    projmatrix_raw: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        return mark_visible(positions, self.raster_settings.viewmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None, theta=None, rho=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        color, depth, opacity, radii, n_touched = _rasterize(
            means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
            self.raster_settings, _lib.VARIANT_B, theta=theta, rho=rho)
        return color, radii, depth, opacity, n_touched
