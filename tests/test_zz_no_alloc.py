"""The allocation-free entry point ``eags_image_l1_forward_static`` (include/eags_raster.h; library-owned static
scratch, no counterpart in the reference) against the existing
``eags_image_l1_forward`` through the C ABI: same sums, same raw gradient images, on two calls in a row with
different inputs (the shared scratch keeps the previous call's contents), within the tolerance of
tests/test_gpu_losses.py (the arithmetic and the reduction order are the same, so the bits are expected equal).

Collected last on purpose (``zz``): a failure here must not hide others.
Runs alone in a fresh process as well as after the rest of the suite."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(H, W, seed, dev):
    g = torch.Generator().manual_seed(seed)
    gt_c = torch.rand(3, H, W, generator=g)
    color = (gt_c + 0.2 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
    gt_d = 0.5 + 4.0 * torch.rand(1, H, W, generator=g)
    gt_d[torch.rand(1, H, W, generator=g) < 0.15] = 0.0
    depth = gt_d + 0.1 * torch.randn(1, H, W, generator=g)
    alpha = torch.rand(1, H, W, generator=g)
    mask = (gt_d[0] > 0).to(torch.uint8)
    return [t.to(dev).contiguous() for t in (color, depth, alpha, gt_c, gt_d, mask)]


def _call(lib, _lib, entry, H, W, tensors, dev, scratch=None):
    color, depth, alpha, gt_c, gt_d, mask = tensors
    f32 = dict(dtype=torch.float32, device=dev)
    # outputs start as garbage that differs between the two entry points: nothing may survive from it
    g_color = torch.full((3, H, W), float("nan"), **f32)
    g_depth = torch.full((1, H, W), float("nan"), **f32)
    g_alpha = torch.full((2, H, W), float("nan"), **f32)
    sums = torch.full((6,), float("nan"), **f32)
    # the tracking loss: alpha^3 weighting, masked colour, normalised by the positive count, eps 1e-8
    args = _lib.EagsImageL1(W, H, 1, 1, 0, 1e-8, color.data_ptr(), depth.data_ptr(), alpha.data_ptr(),
                            gt_c.data_ptr(), gt_d.data_ptr(), mask.data_ptr(), None, 0.0)
    stream = torch.cuda.current_stream(dev).cuda_stream
    torch.cuda.synchronize(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]       # device-wide free memory around the C call alone
    if entry == "static":
        rc = lib.eags_image_l1_forward_static(C.byref(args), g_color.data_ptr(), g_depth.data_ptr(), g_alpha.data_ptr(),
                                              sums.data_ptr(), stream)
    else:
        rc = lib.eags_image_l1_forward(C.byref(args), g_color.data_ptr(), g_depth.data_ptr(), g_alpha.data_ptr(),
                                       scratch.data_ptr(), sums.data_ptr(), stream)
    assert rc == 0, lib.eags_last_error()
    torch.cuda.synchronize(dev)
    return (sums, g_color, g_depth, g_alpha), free_before - torch.cuda.mem_get_info(dev)[0]


def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    from eags_raster import _lib
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    # two calls in a row: different sizes (a ragged one that takes the scalar path, then 640x480 on the vector
    # path) and different data, so stale partial sums or a stale ticket in the shared array would show
    for call, (H, W, seed) in enumerate([(53, 71, 1), (480, 640, 2)]):
        tensors = _inputs(H, W, seed, dev)
        scratch = torch.empty(lib.eags_image_l1_scratch_bytes(W, H), dtype=torch.uint8, device=dev)
        ref, _ = _call(lib, _lib, "alloc", H, W, tensors, dev, scratch)
        got, taken = _call(lib, _lib, "static", H, W, tensors, dev)
        if call == 1:      # (a first call may load the CUDA module lazily; the second one takes no device memory)
            assert taken <= 0, f"the call took {taken} bytes of device memory"
        assert not any(torch.isnan(t).any() for t in got), call
        # tolerance of tests/test_gpu_losses.py: sums 1e-6 relative; the gradient images are signs times weights
        for a, b in zip(got[0].tolist(), ref[0].tolist()):
            assert abs(a - b) <= 1e-6 * max(abs(b), 1e-30), (call, got[0], ref[0])
        for a, b, name in zip(got[1:], ref[1:], ("g_color", "g_depth", "g_alpha")):
            assert float((a - b).abs().max()) <= 1e-6 * float(b.abs().max()), (call, name)
        # same kernel, same grid, fixed-order reduction: the bits are in fact equal
        assert all(torch.equal(a, b) for a, b in zip(got, ref)), call


def test_static_scratch_entry_point_rejects_what_it_cannot_hold():
    from eags_raster import _lib
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    tensors = _inputs(8, 8, 3, dev)
    color, depth, alpha, gt_c, gt_d, mask = tensors
    out = torch.zeros(16, device=dev)
    # 2049 x 2048 pixels > EAGS_IMAGE_L1_STATIC_MAX_PIXELS: refused before anything is launched or read
    args = _lib.EagsImageL1(2049, 2048, 0, 0, 1, 0.0, color.data_ptr(), depth.data_ptr(), None, gt_c.data_ptr(),
                            gt_d.data_ptr(), None, None, 0.0)
    rc = lib.eags_image_l1_forward_static(C.byref(args), out.data_ptr(), out.data_ptr(), None, out.data_ptr(), None)
    assert rc == -1 and b"EAGS_IMAGE_L1_STATIC_MAX_PIXELS" in lib.eags_last_error()
    args = _lib.EagsImageL1(8, 8, 1, 0, 0, 0.0, color.data_ptr(), depth.data_ptr(), None, gt_c.data_ptr(),
                            gt_d.data_ptr(), None, None, 0.0)       # soft_alpha without an alpha image
    rc = lib.eags_image_l1_forward_static(C.byref(args), out.data_ptr(), out.data_ptr(), None, out.data_ptr(), None)
    assert rc == -1 and b"eags_image_l1_forward_static" in lib.eags_last_error()
    torch.cuda.synchronize(dev)
    assert float(out.abs().max()) == 0.0
