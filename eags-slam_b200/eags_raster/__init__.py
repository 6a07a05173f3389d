"""eags_raster -- B200-native (sm_100a) differentiable Gaussian-splatting rasterizer for EAGS-SLAM.

The product path is: ``gaussian_rasterizer`` / ``diff_gaussian_rasterization`` (API shims)
-> :mod:`eags_raster.rasterize` (buffers, streams, autograd) -> ``libeags_raster.so`` (C ABI,
``include/eags_raster.h``) -> hand-written CUDA kernels (``csrc/``).  Nothing here imports ``oracle/``.

Around the operator: :mod:`eags_raster.losses` (fused L1 / SSIM / isotropic losses),
:mod:`eags_raster.tracking` (pose folded into the view matrices), :mod:`eags_raster.neighbors`
(3-NN distances, radius test), :mod:`eags_raster.graph` (CUDA-graph replay of an iteration),
:mod:`eags_raster.peer` / :mod:`eags_raster.sharded` (multi-GPU mapping), :mod:`eags_raster.handoff`.
"""
from . import _lib  # noqa: F401
from .rasterize import (  # noqa: F401
    RasterState,
    get_sync_mode,
    grad_arena,
    rasterize_backward,
    rasterize_forward,
    rasterize_gaussians,
    reserve_capacity,
    set_sync_mode,
    state_array,
)
