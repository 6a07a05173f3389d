"""Drop-in for ``gaussian_rasterizer`` (VladimirYugay/gaussian_rasterizer@9c40173, reference
requirements.txt:21) -- "variant A" of the operator, used by tracking, mapping and evaluation through
``src/utils/utils.py:8,105-117,149,174-184``.

Same names, keyword arguments, output order and error messages as upstream's Python front-end;
the work is done by the sm_100a kernels behind ``libeags_raster.so``.

    color, depth, alpha, radii = GaussianRasterizer(raster_settings=s)(
        means3D=..., means2D=..., opacities=..., shs=... | colors_precomp=...,
        scales=..., rotations=... | cov3D_precomp=...)

``color`` (3,H,W); ``depth`` (1,H,W) = sum_i z_i alpha_i T_i; ``alpha`` (1,H,W) = sum_i alpha_i T_i;
``radii`` (P,) int32.  Gradients flow to means3D, means2D (screen-space), opacities, shs /
colors_precomp, scales, rotations / cov3D_precomp.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from eags_raster import _lib
from eags_raster.rasterize import rasterize_gaussians as _rasterize


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def mark_visible(positions: torch.Tensor, viewmatrix: torch.Tensor) -> torch.Tensor:
    """Frustum visibility mask (upstream ``_C.mark_visible``: view-space z > 0.2).  Never called by
    the reference; provided for API completeness as three torch ops."""
    with torch.no_grad():
        V = viewmatrix.to(positions.dtype)
        z = positions @ V[:3, 2] + V[3, 2]
        return z > 0.2


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        return mark_visible(positions, self.raster_settings.viewmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        color, depth, alpha, radii = _rasterize(means3D, means2D, shs, colors_precomp, opacities,
                                                scales, rotations, cov3D_precomp,
                                                self.raster_settings, _lib.VARIANT_A)
        return color, depth, alpha, radii
