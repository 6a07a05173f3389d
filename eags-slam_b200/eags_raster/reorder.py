"""Storage order of a submap's Gaussians (host-side helper around the binning kernels).

``preprocess_fwd`` reserves bucket space with one global atomic per (CTA, touched tile): a CTA walks a
contiguous span of the Gaussian arrays, so the fewer tiles a span touches the fewer atomics it issues
(csrc/preprocess.cu).  The reference appends new Gaussians keyframe by keyframe from
sampled pixels (src/entities/mapper.py:225-260, gaussian_model.py:586-592), i.e. roughly in image order
already; after pruning / merging (or for a cloud that arrives in arbitrary order) ``tile_order`` gives the
permutation that restores it for a given view.  Applying it changes no rendered value: inside a tile the
blend order is (depth, index) and only exact depth ties -- a measure-zero event for real data -- see the
index.
"""
from __future__ import annotations

from typing import Sequence

import torch


def tile_order(means3D: torch.Tensor, settings) -> torch.Tensor:
    """Permutation (int64, on ``means3D``'s device) that sorts the Gaussians by the 16x16 tile their mean
    projects to under ``settings`` (a GaussianRasterizationSettings: ``viewmatrix`` = w2c^T, utils.py:95-97);
    Gaussians behind the near plane or outside the image go last."""
    V = settings.viewmatrix.to(means3D.device, torch.float32)
    H, W = int(settings.image_height), int(settings.image_width)
    pc = means3D @ V[:3, :3] + V[3, :3]                      # row-vector convention of the stored matrix
    P = settings.projmatrix.to(means3D.device, torch.float32)
    hom = torch.cat([means3D, torch.ones_like(means3D[:, :1])], dim=1) @ P
    ndc = hom[:, :2] / (hom[:, 3:4] + 1e-7)
    u = ((ndc[:, 0] + 1.0) * W - 1.0) * 0.5
    v = ((ndc[:, 1] + 1.0) * H - 1.0) * 0.5
    ui, vi = u.floor().long(), v.floor().long()
    inside = (pc[:, 2] > 0.2) & (ui >= 0) & (ui < W) & (vi >= 0) & (vi < H)
    gx = (W + 15) // 16
    key = torch.where(inside, (vi // 16) * gx + ui // 16, torch.full_like(ui, 1 << 40))
    return torch.argsort(key, stable=True)


def apply_order(perm: torch.Tensor, tensors: Sequence[torch.Tensor]):
    """The per-Gaussian tensors (parameters, optimiser moments, ...) in the new order."""
    return [t[perm].contiguous() for t in tensors]
