"""Fused image-space losses for the tracking / mapping iterations.

Drop-in equivalents of what the reference computes with ~25 small torch kernels per iteration:

* :func:`tracking_losses` -- src/entities/tracker.py:112-128: ``|x - y| + 1e-8`` per element, the
  depth-validity / alpha / outlier ``tracking_mask``, optional ``alpha**3`` weighting, each of the two
  sums divided by its number of positive elements.
* :func:`masked_l1_means` -- the two L1 means of a mapping iteration, src/entities/mapper.py:161-165
  with src/entities/losses.py:8-27 (``l1_loss(image[:, mask], gt[:, mask])``, ``l1_loss(depth[:, mask],
  gt_depth[mask])``); with ``mask=None`` they are plain means (the benchmark's loss).

Both return ``(color_loss, depth_loss)`` as 0-d tensors with autograd to ``color``, ``depth`` (and
``alpha``).  One CUDA kernel produces the four reductions and the raw gradient images, one more scales
them in ``backward`` -- no host synchronisation anywhere.  CUDA tensors only (no CPU path).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib


def _c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.to(torch.float32).contiguous()
    return t


class _ImageL1(torch.autograd.Function):
    @staticmethod
    def forward(ctx, color, depth, alpha, gt_color, gt_depth, mask, soft_alpha, mask_color, norm_mode, eps,
                mask_depth=None, alpha_min=0.0):
        lib = _lib.load()
        if not color.is_cuda:
            raise RuntimeError("eags_raster.losses: tensors must live on a CUDA device (there is no CPU path)")
        dev = color.device
        H, W = int(color.shape[-2]), int(color.shape[-1])
        with torch.cuda.device(dev):
            color, depth, gt_color, gt_depth = _c(color), _c(depth), _c(gt_color), _c(gt_depth)
            if color.numel() != 3 * H * W or depth.numel() != H * W or gt_color.numel() != 3 * H * W or \
                    gt_depth.numel() != H * W:
                raise ValueError("expected color (3,H,W), depth (1,H,W) and matching ground truth")
            alpha_c = _c(alpha) if alpha is not None else None
            if soft_alpha and alpha_c is None:
                raise ValueError("soft_alpha needs the alpha image")
            mask_c = None
            if mask is not None:
                mask_c = mask.reshape(H, W).to(torch.uint8).contiguous()
            mask_d = None
            if mask_depth is not None:
                mask_d = mask_depth.reshape(H, W).to(torch.uint8).contiguous()
            want_alpha_grad = bool(soft_alpha and alpha is not None and ctx.needs_input_grad[2])
            f32 = dict(dtype=torch.float32, device=dev)
            g_color = torch.empty(3, H, W, **f32)
            g_depth = torch.empty(1, H, W, **f32)
            g_alpha = torch.empty(2, H, W, **f32) if want_alpha_grad else None
            sums = torch.empty(6, **f32)
            scratch = torch.empty(lib.eags_image_l1_scratch_bytes(W, H), dtype=torch.uint8, device=dev)
            args = _lib.EagsImageL1(W, H, int(soft_alpha), int(bool(mask_color)), int(norm_mode), float(eps),
                                    color.data_ptr(), depth.data_ptr(),
                                    None if alpha_c is None else alpha_c.data_ptr(), gt_color.data_ptr(),
                                    gt_depth.data_ptr(), None if mask_c is None else mask_c.data_ptr(),
                                    None if mask_d is None else mask_d.data_ptr(), float(alpha_min))
            stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.eags_image_l1_forward(C.byref(args), g_color.data_ptr(), g_depth.data_ptr(),
                                                 None if g_alpha is None else g_alpha.data_ptr(),
                                                 scratch.data_ptr(), sums.data_ptr(), stream),
                       "eags_image_l1_forward")
            # loss = sum / normaliser, formed by the kernel (0/0 -> nan exactly as the reference's division)
            color_loss, depth_loss = sums[4], sums[5]
        ctx.save_for_backward(sums, g_color, g_depth, g_alpha if g_alpha is not None else torch.empty(0))
        ctx.shape = (H, W)
        ctx.alpha_shape = None if alpha is None else alpha.shape
        ctx.shapes = (color.shape, depth.shape)
        return color_loss, depth_loss

    @staticmethod
    def backward(ctx, up_color, up_depth):
        lib = _lib.load()
        sums, g_color, g_depth, g_alpha = ctx.saved_tensors
        H, W = ctx.shape
        dev = sums.device
        with torch.cuda.device(dev):
            zero = None
            if up_color is None or up_depth is None:
                zero = torch.zeros((), dtype=torch.float32, device=dev)
            up_color = _c(up_color) if up_color is not None else zero
            up_depth = _c(up_depth) if up_depth is not None else zero
            ga = g_alpha if g_alpha.numel() else None
            stream = torch.cuda.current_stream(dev).cuda_stream
            # the raw gradients are scaled in place: a second backward through the same node is not supported
            _lib.check(lib.eags_image_l1_backward(W, H, sums.data_ptr(), up_color.data_ptr(), up_depth.data_ptr(),
                                                  g_color.data_ptr(), g_depth.data_ptr(),
                                                  None if ga is None else ga.data_ptr(), stream),
                       "eags_image_l1_backward")
        d_alpha = None
        if ga is not None:
            d_alpha = ga[0].view(ctx.alpha_shape)
        return (g_color.view(ctx.shapes[0]), g_depth.view(ctx.shapes[1]), d_alpha, None, None, None, None, None,
                None, None, None, None)


def tracking_losses(color: torch.Tensor, depth: torch.Tensor, alpha: torch.Tensor, gt_color: torch.Tensor,
                    gt_depth: torch.Tensor, tracking_mask: Optional[torch.Tensor], soft_alpha: bool = True,
                    mask_invalid_depth_in_color_loss: bool = False, eps: float = 1e-8
                    ) -> Tuple[torch.Tensor, torch.Tensor]:
    """``color_loss, depth_loss`` of reference tracker.py:116-128 (``tracking_mask`` as built at :102-113)."""
    return _ImageL1.apply(color, depth, alpha, gt_color, gt_depth, tracking_mask, bool(soft_alpha),
                          bool(mask_invalid_depth_in_color_loss), 0, float(eps))


def masked_l1_means(color: torch.Tensor, depth: torch.Tensor, gt_color: torch.Tensor, gt_depth: torch.Tensor,
                    mask: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``l1_loss(color[:, mask], gt_color[:, mask]), l1_loss(depth[:, mask], gt_depth[mask])`` of reference
    mapper.py:161-165; plain means when ``mask`` is None."""
    return _ImageL1.apply(color, depth, None, gt_color, gt_depth, mask, False, False, 1, 0.0)


def registration_loss(image: torch.Tensor, depth: torch.Tensor, opacity: torch.Tensor, gt_image: torch.Tensor,
                      gt_depth: torch.Tensor, rgb_mask: Optional[torch.Tensor] = None,
                      depth_mask: Optional[torch.Tensor] = None, alpha: float = 0.95,
                      opacity_min: float = 0.95) -> torch.Tensor:
    """``get_loss_tracking_rgbd`` of the loop-closure registration (reference src/gsr/loss.py:66-95), two kernels:
    ``alpha * mean(opacity * |image*m - gt*m|) + (1 - alpha) * mean(|depth*dm - gt_depth*dm|)`` with
    ``m = rgb_mask`` (``gt.sum(0) > rgb_boundary_threshold`` times the camera's grad mask, loss.py:71-75) and
    ``dm = depth_mask * (opacity > 0.95)`` (``depth_mask = gt_depth > 0.01``, loss.py:88-92).  Masks are
    (H,W) / (1,H,W) bool or uint8; None = all valid.  Gradients flow to image, depth and opacity."""
    l_rgb, l_depth = _ImageL1.apply(image, depth, opacity, gt_image, gt_depth, rgb_mask, 2, True, 2, 0.0,
                                    depth_mask, float(opacity_min))
    return alpha * l_rgb + (1.0 - alpha) * l_depth


class _Ssim(torch.autograd.Function):
    """mean SSIM(img1, img2), gradient to ``img1`` (the rendered image); ``img2`` is the ground truth."""

    @staticmethod
    def forward(ctx, img1, img2):
        lib = _lib.load()
        if not img1.is_cuda:
            raise RuntimeError("eags_raster.losses: tensors must live on a CUDA device (there is no CPU path)")
        if img1.shape != img2.shape or img1.dim() not in (3, 4) or (img1.dim() == 4 and img1.shape[0] != 1):
            raise ValueError("ssim expects two (C,H,W) or (1,C,H,W) images of the same shape")
        dev = img1.device
        Cc, H, W = (int(v) for v in img1.shape[-3:])
        with torch.cuda.device(dev):
            a, b = _c(img1), _c(img2)
            scratch = torch.empty(lib.eags_ssim_scratch_bytes(Cc, H, W), dtype=torch.uint8, device=dev)
            out = torch.empty((), dtype=torch.float32, device=dev)
            stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.eags_ssim_forward(Cc, H, W, a.data_ptr(), b.data_ptr(), scratch.data_ptr(), out.data_ptr(),
                                             stream), "eags_ssim_forward")
        ctx.save_for_backward(a, b, scratch)
        ctx.dims = (Cc, H, W)
        ctx.in_shape = img1.shape
        return out

    @staticmethod
    def backward(ctx, up):
        lib = _lib.load()
        a, b, scratch = ctx.saved_tensors
        Cc, H, W = ctx.dims
        dev = a.device
        with torch.cuda.device(dev):
            up = _c(up).reshape(())
            grad = torch.empty(Cc, H, W, dtype=torch.float32, device=dev)
            stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.eags_ssim_backward(Cc, H, W, a.data_ptr(), b.data_ptr(), scratch.data_ptr(), up.data_ptr(),
                                              grad.data_ptr(), stream), "eags_ssim_backward")
        return grad.view(ctx.in_shape), None


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11, size_average: bool = True) -> torch.Tensor:
    """``ssim(img1, img2)`` of reference src/entities/losses.py:65-85 (same signature).  The fused kernels
    implement what the reference calls: the 11x11 window with ``size_average=True``; gradient flows to
    ``img1``."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("the fused SSIM implements window_size=11, size_average=True (mapper.py:163)")
    if img2.requires_grad:
        raise NotImplementedError("the fused SSIM differentiates with respect to img1 only (img2 is the ground truth)")
    return _Ssim.apply(img1, img2)


def mapping_color_loss(image: torch.Tensor, gt_image: torch.Tensor, mask: Optional[torch.Tensor],
                       lambda_dssim: float) -> torch.Tensor:
    """``(1 - lambda) * l1_loss(image[:, mask], gt[:, mask]) + lambda * (1 - ssim(image, gt))`` of reference
    mapper.py:162-163, from the fused L1 and SSIM kernels."""
    depth_dummy = image[:1]
    l1, _ = _ImageL1.apply(image, depth_dummy, None, gt_image, depth_dummy.detach(), mask, False, True, 1, 0.0)
    return (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - ssim(image, gt_image))


class _MappingLoss(torch.autograd.Function):
    """(colour loss, depth loss) of one mapping view; one forward and one backward kernel."""

    @staticmethod
    def forward(ctx, image, depth, gt_image, gt_depth, mask, lambda_dssim):
        lib = _lib.load()
        if not image.is_cuda:
            raise RuntimeError("eags_raster.losses: tensors must live on a CUDA device (there is no CPU path)")
        dev = image.device
        H, W = int(image.shape[-2]), int(image.shape[-1])
        with torch.cuda.device(dev):
            image, depth, gt_image, gt_depth = _c(image), _c(depth), _c(gt_image), _c(gt_depth)
            if image.numel() != 3 * H * W or depth.numel() != H * W or gt_image.numel() != 3 * H * W or \
                    gt_depth.numel() != H * W:
                raise ValueError("expected image (3,H,W), depth (1,H,W) and matching ground truth")
            valid_depth = isinstance(mask, str)
            if valid_depth and mask != "valid_depth":
                raise ValueError("mapping_losses: mask is a tensor, None or the string 'valid_depth'")
            mask_c = None if (mask is None or valid_depth) else mask.reshape(H, W).to(torch.uint8).contiguous()
            out = torch.empty(8, dtype=torch.float32, device=dev)
            scratch = torch.empty(lib.eags_mapping_loss_scratch_bytes(W, H), dtype=torch.uint8, device=dev)
            args = _lib.EagsMappingLoss(W, H, float(lambda_dssim), image.data_ptr(), depth.data_ptr(),
                                        gt_image.data_ptr(), gt_depth.data_ptr(),
                                        None if mask_c is None else mask_c.data_ptr(), int(valid_depth))
            _lib.check(lib.eags_mapping_loss_forward(C.byref(args), scratch.data_ptr(), out.data_ptr(),
                                                     torch.cuda.current_stream(dev).cuda_stream),
                       "eags_mapping_loss_forward")
        ctx.save_for_backward(image, depth, gt_image, gt_depth, scratch, out,
                              mask_c if mask_c is not None else torch.empty(0))
        ctx.cfg = (H, W, float(lambda_dssim), int(valid_depth))
        ctx.shapes = (image.shape, depth.shape)
        return out[7], out[5]

    @staticmethod
    def backward(ctx, up_color, up_depth):
        lib = _lib.load()
        image, depth, gt_image, gt_depth, scratch, out, mask_c = ctx.saved_tensors
        H, W, lam, valid_depth = ctx.cfg
        dev = image.device
        with torch.cuda.device(dev):
            zero = None
            if up_color is None or up_depth is None:
                zero = torch.zeros((), dtype=torch.float32, device=dev)
            up_color = _c(up_color) if up_color is not None else zero
            up_depth = _c(up_depth) if up_depth is not None else zero
            g_image = torch.empty(3, H, W, dtype=torch.float32, device=dev)
            g_depth = torch.empty(1, H, W, dtype=torch.float32, device=dev)
            args = _lib.EagsMappingLoss(W, H, lam, image.data_ptr(), depth.data_ptr(), gt_image.data_ptr(),
                                        gt_depth.data_ptr(), mask_c.data_ptr() if mask_c.numel() else None,
                                        valid_depth)
            _lib.check(lib.eags_mapping_loss_backward(C.byref(args), scratch.data_ptr(), out.data_ptr(),
                                                      up_color.data_ptr(), up_depth.data_ptr(), g_image.data_ptr(),
                                                      g_depth.data_ptr(),
                                                      torch.cuda.current_stream(dev).cuda_stream),
                       "eags_mapping_loss_backward")
        return g_image.view(ctx.shapes[0]), g_depth.view(ctx.shapes[1]), None, None, None, None


def mapping_losses(image: torch.Tensor, depth: torch.Tensor, gt_image: torch.Tensor, gt_depth: torch.Tensor,
                   mask: Optional[torch.Tensor], lambda_dssim: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(color_loss, depth_loss)`` of one mapping view, reference src/entities/mapper.py:161-165::

        color_loss = (1 - lambda) * l1_loss(image[:, mask], gt[:, mask]) + lambda * (1 - ssim(image, gt))
        depth_loss = l1_loss(depth[:, mask], gt_depth[mask])

    in ONE forward and ONE backward kernel (the masked L1 terms ride on the SSIM passes; csrc/ssim.cu).
    ``mask``: (H,W) / (1,H,W) bool or uint8, None = all valid, or the string ``"valid_depth"``: the mapper's own
    mask ``(gt_depth > 0) & ~isnan(depth)`` (mapper.py:159) formed per pixel inside the kernels, without the
    three torch kernels and the uint8 copy a mask tensor costs per view.  Gradients flow to ``image`` and ``depth``.
    The same values as ``mapping_color_loss`` + ``masked_l1_means`` (which stay for callers that want the
    pieces), without their raw-gradient images, scaling passes and autograd adds."""
    return _MappingLoss.apply(image, depth, gt_image, gt_depth, mask, lambda_dssim)


class _Isotropic(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scaling):
        lib = _lib.load()
        if not scaling.is_cuda:
            raise RuntimeError("eags_raster.losses: tensors must live on a CUDA device (there is no CPU path)")
        if scaling.dim() != 2 or scaling.shape[1] != 3 or scaling.shape[0] == 0:
            raise ValueError("isotropic_loss expects a non-empty (P, 3) scaling tensor")
        dev = scaling.device
        with torch.cuda.device(dev):
            s = _c(scaling)
            P = int(s.shape[0])
            grad = torch.empty_like(s)
            acc = torch.empty(1, dtype=torch.float64, device=dev)
            out = torch.empty((), dtype=torch.float32, device=dev)
            _lib.check(lib.eags_isotropic_loss(P, s.data_ptr(), grad.data_ptr(), acc.data_ptr(), out.data_ptr(),
                                               torch.cuda.current_stream(dev).cuda_stream), "eags_isotropic_loss")
        ctx.save_for_backward(grad)
        ctx.inv = 1.0 / (3.0 * P)
        return out

    @staticmethod
    def backward(ctx, up):
        (grad,) = ctx.saved_tensors
        return grad * (up * ctx.inv)


def isotropic_loss(scaling: torch.Tensor) -> torch.Tensor:
    """``isotropic_loss(scaling)`` of reference src/entities/losses.py:115-126 (mapper.py:166), one kernel."""
    return _Isotropic.apply(scaling)
