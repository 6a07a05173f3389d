"""Tracking fast path: render under an optimisable relative camera pose without touching the
Gaussians.

The reference's tracking iteration (src/entities/tracker.py:85-97) builds ``rel_transform`` from a
quaternion + translation, moves ALL P means and left-multiplies ALL P rotations in PyTorch, renders
them with ``render_gaussian_model(override_means_3d=..., override_rotations=...)`` from the fixed
previous camera, and gets the pose gradient by autograd through ``dL/dmeans3D`` (P,3) and
``dL/drotations`` (P,4): ~40 P-sized torch kernels per iteration.  Rendering the moved Gaussians
from camera ``L`` equals rendering the original ones from camera ``L @ Rel``, so here

    color, depth, alpha, radii = render_relative(settings, means3D, opacities, shs, scales, rotations,
                                                 cam_rot, cam_trans)

folds the pose into the view matrices (one tiny kernel), renders with the ordinary kernels, and in
backward turns the rasterizer's own camera-pose gradient ``dL/dtau`` into ``cam_rot.grad`` (4,) and
``cam_trans.grad`` (3,) (one tiny kernel).  The per-Gaussian gradient arrays are never written.

Same image as the reference path up to fp32 rounding (the pose is applied once to the camera instead
of P times to the points).  SH degree must be 0 -- what every call site of the reference uses
(utils.py:77,114): the reference evaluates view-dependent SH in the *moved* frame, which this
formulation does not reproduce.
"""
from __future__ import annotations

import ctypes as C
from types import SimpleNamespace

import torch

from . import _lib
from .rasterize import _f32c, rasterize_backward, rasterize_forward


def compose_settings(settings, cam_rot: torch.Tensor, cam_trans: torch.Tensor):
    """Settings of the camera ``L @ Rel(cam_rot, cam_trans)``; also returns the contiguous originals."""
    lib = _lib.load()
    device = cam_rot.device
    with torch.cuda.device(device):
        vm = _f32c(settings.viewmatrix, device)
        pm = _f32c(settings.projmatrix, device)
        cp = _f32c(settings.campos, device)
        q = _f32c(cam_rot.detach().reshape(-1), device)
        t = _f32c(cam_trans.detach().reshape(-1), device)
        if q.numel() != 4 or t.numel() != 3:
            raise ValueError("cam_rot must have 4 elements (wxyz) and cam_trans 3")
        out = torch.empty(35, dtype=torch.float32, device=device)
        stream = torch.cuda.current_stream(device).cuda_stream
        _lib.check(lib.eags_pose_compose(q.data_ptr(), t.data_ptr(), vm.data_ptr(), pm.data_ptr(), cp.data_ptr(),
                                         out.data_ptr(), out.data_ptr() + 64, out.data_ptr() + 128, stream),
                   "eags_pose_compose")
    s2 = SimpleNamespace(
        image_height=settings.image_height, image_width=settings.image_width, tanfovx=settings.tanfovx,
        tanfovy=settings.tanfovy, bg=settings.bg, scale_modifier=settings.scale_modifier,
        viewmatrix=out[:16].view(4, 4), projmatrix=out[16:32].view(4, 4), sh_degree=settings.sh_degree,
        campos=out[32:35], prefiltered=settings.prefiltered, debug=settings.debug,
        projmatrix_raw=getattr(settings, "projmatrix_raw", None))
    return s2, (q, t, vm)


class _RenderRelative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cam_rot, cam_trans, means3D, opacities, shs, scales, rotations, settings):
        if not means3D.is_cuda:
            raise RuntimeError("eags_raster.tracking: tensors must live on a CUDA device (there is no CPU path)")
        if int(settings.sh_degree) != 0:
            raise ValueError("render_relative supports sh_degree == 0 only (see the module docstring)")
        s2, (q, t, vm) = compose_settings(settings, cam_rot, cam_trans)
        color, depth, alpha, radii, _, state = rasterize_forward(
            s2, _lib.VARIANT_A, means3D, opacities, shs=shs, scales=scales, rotations=rotations)
        ctx.s2, ctx.state, ctx.pose = s2, state, (q, t, vm)
        ctx.shapes = (cam_rot.shape, cam_trans.shape)
        ctx.save_for_backward(means3D, opacities, shs, scales, rotations)
        ctx.mark_non_differentiable(radii)
        return color, depth, alpha, radii

    @staticmethod
    def backward(ctx, g_color, g_depth, g_alpha, _unused):
        lib = _lib.load()
        means3D, opacities, shs, scales, rotations = ctx.saved_tensors
        q, t, vm = ctx.pose
        device = q.device
        if ctx.state.P == 0:
            return (torch.zeros(ctx.shapes[0], device=device), torch.zeros(ctx.shapes[1], device=device),
                    None, None, None, None, None, None)
        out = rasterize_backward(ctx.s2, ctx.state, means3D, opacities, shs, None, scales, rotations, None,
                                 g_color, g_depth, g_alpha, want_shs=False, want_tau=True, pose_only=True)
        with torch.cuda.device(device):
            g = torch.empty(7, dtype=torch.float32, device=device)
            stream = torch.cuda.current_stream(device).cuda_stream
            _lib.check(lib.eags_pose_chain(q.data_ptr(), t.data_ptr(), vm.data_ptr(), out["tau"].data_ptr(),
                                           g.data_ptr(), g.data_ptr() + 16, stream), "eags_pose_chain")
        return g[:4].view(ctx.shapes[0]), g[4:].view(ctx.shapes[1]), None, None, None, None, None, None


def render_relative(settings, means3D, opacities, shs, scales, rotations, cam_rot, cam_trans):
    """Variant-A render of (means3D, ..., rotations) moved by the relative pose (cam_rot wxyz, un-normalised
    as ``opt_cam_rot`` in tracker.py:85; cam_trans) seen from ``settings``' camera.  Returns
    ``(color, depth, alpha, radii)``; gradients flow to ``cam_rot`` and ``cam_trans`` only."""
    return _RenderRelative.apply(cam_rot, cam_trans, means3D, opacities, shs, scales, rotations, settings)


def update_pose_se3(theta: torch.Tensor, rho: torch.Tensor, viewmatrix: torch.Tensor, projmatrix_raw: torch.Tensor,
                    projmatrix: torch.Tensor, campos: torch.Tensor, tau_norm: torch.Tensor = None) -> None:
    """The per-iteration camera update of loop-closure registration, on the device and in place: the
    reference's ``update_pose(viewpoint)`` (src/utils/pose_utils.py:76-93: ``w2c <- SE3_exp([rho, theta]) @ w2c``,
    deltas zeroed) plus the recomputation of ``world_view_transform`` / ``full_proj_transform`` /
    ``camera_center`` (src/gsr/camera.py:97-111).  ``viewmatrix`` (= w2c^T), ``projmatrix`` and ``campos`` are
    the tensors the rasterizer settings hold: they are overwritten, so a CUDA-graph replay of the iteration sees
    the new camera.  ``tau_norm`` (1-element device tensor) receives ``|tau|`` -- the reference's convergence
    test reads it (``< 1e-4``) without forcing a sync per iteration."""
    lib = _lib.load()
    dev = theta.device
    for name, x, n in (("theta", theta, 3), ("rho", rho, 3), ("viewmatrix", viewmatrix, 16),
                       ("projmatrix_raw", projmatrix_raw, 16), ("projmatrix", projmatrix, 16), ("campos", campos, 3)):
        if not (x.is_cuda and x.device == dev and x.dtype == torch.float32 and x.is_contiguous() and x.numel() == n):
            raise ValueError(f"update_pose_se3: {name} must be a contiguous fp32 CUDA tensor of {n} elements on {dev}")
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(lib.eags_pose_se3_update(theta.data_ptr(), rho.data_ptr(), viewmatrix.data_ptr(),
                                            projmatrix_raw.data_ptr(), projmatrix.data_ptr(), campos.data_ptr(),
                                            None if tau_norm is None else tau_norm.data_ptr(), stream),
                   "eags_pose_se3_update")


class PoseAdamSE3:
    """The optimiser side of one loop-closure registration iteration as ONE kernel
    (``eags_pose_adam_se3_update``): ``optimizer.step()`` on the two pose deltas (torch.optim.Adam, lr 3 x base
    for ``cam_rot_delta``, base for ``cam_trans_delta``; reference src/gsr/solver.py:38-51,83-91) followed by
    ``update_pose`` (pose_utils.py:76-93) and the camera properties (camera.py:97-111).  Replaces ~15 tiny
    optimiser kernels per iteration; CUDA-graph friendly (the step count lives on the device)."""

    def __init__(self, device, base_lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        self.state = torch.zeros(13, dtype=torch.float32, device=device)
        self.lr_rot, self.lr_trans, self.betas, self.eps = 3.0 * base_lr, base_lr, betas, eps

    def step(self, theta: torch.Tensor, rho: torch.Tensor, viewmatrix, projmatrix_raw, projmatrix, campos,
             tau_norm=None) -> None:
        """Uses ``theta.grad`` / ``rho.grad``; the deltas themselves stay zero (as after ``update_pose``)."""
        lib = _lib.load()
        dev = self.state.device
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.eags_pose_adam_se3_update(
                theta.grad.data_ptr(), rho.grad.data_ptr(), self.state.data_ptr(), self.lr_rot, self.lr_trans,
                self.betas[0], self.betas[1], self.eps, viewmatrix.data_ptr(), projmatrix_raw.data_ptr(),
                projmatrix.data_ptr(), campos.data_ptr(), None if tau_norm is None else tau_norm.data_ptr(), stream),
                "eags_pose_adam_se3_update")
