"""Host side of the rasterizer: state buffers, stream plumbing and the autograd bridge.

Mirrors what upstream's ``diff_gaussian_rasterization/__init__.py::_RasterizeGaussians`` does around
its two pybind calls, but over the C ABI of ``include/eags_raster.h``:
PyTorch allocates every buffer (geom / binning / image state, outputs, gradients) and hands raw
device pointers plus the current stream to the library.

Host synchronisation.  Upstream blocks on a device->host copy of ``num_rendered`` in the middle of
every forward to size the binning buffers.  Here the forward is issued *optimistically*: stage 1
(``eags_forward_geom``) enqueues an async copy of K into pinned memory, stage 2
(``eags_forward_render``) is enqueued immediately against a binning buffer whose capacity is a
running high-water mark, and only then does the host wait -- for the stage-1 event, not for the
render -- so the GPU never idles.  If K exceeded the capacity the render kernels exited without
touching memory and stage 2 is re-issued with a larger buffer.  ``set_sync_mode("exact")``
reproduces upstream's behaviour (wait, then allocate exactly K); ``"nosync"`` never waits (for CUDA
graph capture; an overflow raises on the next call).
"""
from __future__ import annotations

import ctypes as C
import os
import threading
import weakref
from typing import NamedTuple, Optional

import torch

from . import _lib
from ._lib import EagsGaussians, EagsImageGrads, EagsImages, EagsInputGrads, EagsView

_SYNC_MODES = ("optimistic", "exact", "nosync")
_sync_mode = "optimistic"
_MIN_CAPACITY = 1 << 16
_capacity = {}            # device index -> high-water capacity (keys)
_tls = threading.local()  # pinned K slots / pending overflow checks per thread
_tile_cap = {}            # (device index, W, H) -> bucket slots per tile for the bucketed forward
_bucketed = os.environ.get("EAGS_BUCKETED", "1") != "0"   # optimistic mode: scatter keys in stage 1
_MAX_BUCKET_SLOTS = 1 << 28                               # 2 GiB of keys; beyond that use the exact path


_kernel_stats = False


def enable_kernel_stats(on: bool) -> None:
    """Measurement only (bench.py's executed-work roofline): the blend kernels count the (8x4 block, splat)
    iterations they walk after the cull and the (pixel, splat) pairs they blend.  Read them with
    :func:`forward_stats` and from ``rasterize_backward(...)["stats"]``."""
    global _kernel_stats
    _kernel_stats = bool(on)
    _lib.set_option("kernel_stats", int(on))


def forward_stats(state) -> dict:
    ctrl = state_array(state, "ctrl").view(torch.int64)
    return {"warp_iters": int(ctrl[5]), "lane_pairs": int(ctrl[6])}


def set_bucketed(on: bool) -> None:
    """Optimistic mode only: let stage 1 scatter the keys into per-tile buckets sized from the
    previous frames' largest tile (no duplicate kernel); off = always the exact path."""
    global _bucketed
    _bucketed = bool(on)


def set_sync_mode(mode: str) -> None:
    global _sync_mode
    if mode not in _SYNC_MODES:
        raise ValueError(f"sync mode must be one of {_SYNC_MODES}")
    _sync_mode = mode


def get_sync_mode() -> str:
    return _sync_mode


def reserve_capacity(device, num_keys: int) -> None:
    """Pre-size the binning buffer (e.g. before ``nosync`` / graph capture)."""
    idx = torch.device(device).index
    idx = torch.cuda.current_device() if idx is None else idx
    _capacity[idx] = max(_capacity.get(idx, 0), int(num_keys))


# Keyed by DEVICE INDEX, not process-wide and not thread-local: autograd runs backward() on its own per-device
# thread, so a thread-local would never be seen by the backward it is meant for, while a process-wide value
# would leak into the loop-closure thread that renders on another device (reference src/entities/lc.py:625)
# while sharded mapping runs on this one.  (A pose-only backward -- registration -- ignores both anyway.)
_arena = {}          # device index -> dict of target tensors (grad_arena)
_peer_route = {}     # device index -> _RouteCtx (eags_raster.peer.PeerGradExchange.route())
_capture_ctx = None  # set by eags_raster.graph.GraphedIteration while it captures (capture is process-exclusive)


class grad_arena:
    """Context manager: while active, the backward pass writes the gradients of the named operator
    inputs straight into caller-provided tensors (e.g. slices of one fused all-reduce buffer,
    :class:`eags_raster.sharded.FusedGradBucket`) instead of fresh allocations, so a following
    collective needs no packing.  Valid for ONE backward per activation (a second backward would
    overwrite the first result).  Applies to backward passes on the DEVICE of the given tensors only, so a
    second thread rendering on another device (the loop-closure worker) is unaffected; keys: means3D,
    scales, rotations, opacities, shs, colors_precomp, cov3D_precomp, means2D."""

    def __init__(self, **tensors):
        self.tensors = tensors
        devs = {t.device.index for t in tensors.values() if t is not None and t.is_cuda}
        if len(devs) != 1:
            raise ValueError("grad_arena: the target tensors must live on one CUDA device")
        self.index = devs.pop()

    def __enter__(self):
        self.prev = _arena.get(self.index)
        _arena[self.index] = self.tensors
        return self

    def __exit__(self, *exc):
        if self.prev is None:
            _arena.pop(self.index, None)
        else:
            _arena[self.index] = self.prev
        return False


class RasterState(NamedTuple):
    """Everything the backward pass (and the parity tests) need from a forward pass."""
    P: int
    M: int
    W: int
    H: int
    variant: int
    capacity: int
    with_keys: bool
    num_rendered: int          # -1 in nosync mode (not known on the host)
    radii: torch.Tensor
    geom: torch.Tensor
    image: torch.Tensor
    binning: torch.Tensor


# ---- small helpers -------------------------------------------------------------------------------
_cont_cache = {}


def _f32c(t: torch.Tensor, device) -> torch.Tensor:
    """fp32 + contiguous + on ``device``.  The reference passes ``w2c.transpose(0, 1)`` views
    (utils.py:97) every call; the materialised copy is cached per (tensor, version)."""
    if t.dtype == torch.float32 and t.is_contiguous() and t.device == device:
        return t
    if torch.cuda.is_current_stream_capturing():
        # CUDA-graph capture: the copy must be a node of the graph (re-done from the source view on every
        # replay, allocated in the graph's pool) -- never a cached tensor that lives outside it
        return t.detach().to(device=device, dtype=torch.float32).contiguous()
    key = id(t)
    hit = _cont_cache.get(key)
    if hit is not None:
        ref, ver, out = hit
        if ref() is t and ver == t._version:
            return out
    out = t.detach().to(device=device, dtype=torch.float32).contiguous()
    if len(_cont_cache) > 256:
        # drop entries whose source tensor is gone; never the whole cache (another thread's kernels may
        # still be reading a live entry's copy)
        for k in [k for k, (ref, _, _) in list(_cont_cache.items()) if ref() is None]:
            _cont_cache.pop(k, None)
        while len(_cont_cache) > 1024:
            _cont_cache.pop(next(iter(_cont_cache)), None)
    try:
        _cont_cache[key] = (weakref.ref(t), t._version, out)
    except TypeError:
        pass
    return out


_size_cache = {}


def _sizes(lib, name: str, *args) -> int:
    """Memoised eags_*_bytes queries (pure functions of their arguments)."""
    key = (name,) + args
    v = _size_cache.get(key)
    if v is None:
        if len(_size_cache) > 4096:
            _size_cache.clear()
        v = _size_cache[key] = int(getattr(lib, name)(*args))
    return v


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _pinned_slot() -> torch.Tensor:
    slot = getattr(_tls, "slot", None)
    if slot is None:
        slot = torch.zeros(1, dtype=torch.int64).pin_memory()
        _tls.slot = slot
    return slot


def _nosync_slot() -> torch.Tensor:
    """nosync mode: a pinned K slot from a per-thread ring.  cudaHostAlloc synchronises the device, so slots
    are allocated once and recycled as soon as the forward that owned them is no longer pending."""
    ring = getattr(_tls, "ring", None)
    if ring is None:
        ring = _tls.ring = []
    busy = {id(p[2]) for p in (getattr(_tls, "pending", None) or [])}
    for s in ring:
        if id(s) not in busy:
            return s
    s = torch.zeros(1, dtype=torch.int64).pin_memory()
    ring.append(s)
    return s


def _pinned_status() -> torch.Tensor:
    st = getattr(_tls, "status", None)
    if st is None:
        st = torch.zeros(8, dtype=torch.int64).pin_memory()   # EagsForwardStatus, 64 bytes
        _tls.status = st
    return st


def _round_capacity(k: int) -> int:
    k = max(int(k), _MIN_CAPACITY)
    return (k + 0xFFFF) & ~0xFFFF


_view_cache = {}   # (id(settings), variant, device index) -> (settings, EagsView, keep)


def _make_view(s, variant: int, device) -> tuple:
    """EagsView of a settings tuple.  The struct holds POINTERS, so when every camera tensor is already a
    contiguous fp32 tensor on the device (values may change in place, addresses do not) it is built once per
    settings object and reused -- the tracking / mapping loops call with the same settings 100+ times."""
    key = (id(s), variant, device.index)
    hit = _view_cache.get(key)
    if hit is not None and hit[0] is s:
        return hit[1], hit[2]
    v, keep, direct = _build_view(s, variant, device)
    if direct and isinstance(s, tuple):      # NamedTuple settings are immutable; duck-typed objects are not cached
        if len(_view_cache) > 256:
            _view_cache.clear()          # plain host structs: nothing on the device refers to them
        _view_cache[key] = (s, v, keep)
    return v, keep


def _build_view(s, variant: int, device) -> tuple:
    bg = _f32c(s.bg, device)
    vm = _f32c(s.viewmatrix, device)
    pm = _f32c(s.projmatrix, device)
    cp = _f32c(s.campos, device)
    raw = getattr(s, "projmatrix_raw", None)
    raw = None if raw is None else _f32c(raw, device)
    v = EagsView(int(s.image_height), int(s.image_width), float(s.tanfovx), float(s.tanfovy),
                 float(s.scale_modifier), int(s.sh_degree), int(variant), int(bool(s.prefiltered)),
                 int(bool(s.debug)), bg.data_ptr(), vm.data_ptr(), pm.data_ptr(), _ptr(raw),
                 cp.data_ptr())
    srcs = (s.bg, s.viewmatrix, s.projmatrix, s.campos, getattr(s, "projmatrix_raw", None))
    direct = all(a is b for a, b in zip(srcs, (bg, vm, pm, cp, raw)))
    return v, (bg, vm, pm, cp, raw), direct   # keep the tensors alive while the struct is in use


def _make_gaussians(P, M, means3D, opacities, shs, colors_precomp, scales, rotations, cov3D):
    return EagsGaussians(P, M, _ptr(means3D), _ptr(opacities), _ptr(shs), _ptr(colors_precomp),
                         _ptr(scales), _ptr(rotations), _ptr(cov3D))


def _prep(t: Optional[torch.Tensor], device) -> Optional[torch.Tensor]:
    if t is None or t.numel() == 0:
        return None
    if t.dtype != torch.float32 or not t.is_contiguous() or t.device != device:
        t = t.to(device=device, dtype=torch.float32).contiguous()
    return t


def _check_pending(idx: int) -> None:
    pend = getattr(_tls, "pending", None)
    if not pend:
        return
    keep = []
    for (dev, ev, slot, cap) in pend:
        if ev.query():
            k = int(slot[0])
            if k > cap:
                _capacity[dev] = _round_capacity(int(k * 1.5))
                _tls.pending = []
                raise RuntimeError(
                    f"eags_raster (nosync mode): a previous forward produced {k} keys but the binning "
                    f"buffer held {cap}; its outputs were not rendered. Capacity has been raised.")
        else:
            keep.append((dev, ev, slot, cap))
    _tls.pending = keep


# ---- forward / backward over the C ABI ------------------------------------------------------------
def rasterize_forward(settings, variant, means3D, opacities, shs=None, colors_precomp=None,
                      scales=None, rotations=None, cov3D_precomp=None, want_n_touched=False,
                      with_keys=False):
    """Returns (color (3,H,W), depth (1,H,W), third (1,H,W), radii (P,), n_touched | None, state)."""
    lib = _lib.load()
    if not means3D.is_cuda:
        raise RuntimeError("eags_raster: tensors must live on a CUDA device (there is no CPU path)")
    device = means3D.device
    H, W = int(settings.image_height), int(settings.image_width)
    with torch.cuda.device(device):
        means3D = _prep(means3D, device)
        P = 0 if means3D is None else int(means3D.shape[0])
        opacities = _prep(opacities, device)
        shs = _prep(shs, device)
        colors_precomp = _prep(colors_precomp, device)
        scales = _prep(scales, device)
        rotations = _prep(rotations, device)
        cov3D_precomp = _prep(cov3D_precomp, device)
        if P > 0:
            if (shs is None) == (colors_precomp is None):
                raise ValueError("Please provide exactly one of either SHs or precomputed colors!")
            if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                    ((scales is not None or rotations is not None) and cov3D_precomp is not None):
                raise ValueError("Please provide exactly one of either scale/rotation pair or "
                                 "precomputed 3D covariance!")
        M = int(shs.shape[1]) if shs is not None else 0
        stream = torch.cuda.current_stream(device).cuda_stream
        view, keep = _make_view(settings, variant, device)
        g = _make_gaussians(P, M, means3D, opacities, shs, colors_precomp, scales, rotations,
                            cov3D_precomp)
        u8 = dict(dtype=torch.uint8, device=device)
        geom = torch.empty(_sizes(lib, "eags_geom_state_bytes", P), **u8)
        image = torch.empty(_sizes(lib, "eags_image_state_bytes", W, H), **u8)
        radii = torch.empty(P, dtype=torch.int32, device=device)
        color = torch.empty(3, H, W, dtype=torch.float32, device=device)
        depth = torch.empty(1, H, W, dtype=torch.float32, device=device)
        third = torch.empty(1, H, W, dtype=torch.float32, device=device)
        n_touched = torch.empty(P, dtype=torch.int32, device=device) if want_n_touched else None
        imgs = EagsImages(color.data_ptr(), depth.data_ptr(), third.data_ptr(), _ptr(n_touched))
        idx = device.index
        mode = _sync_mode
        capturing = P > 0 and torch.cuda.is_current_stream_capturing()
        if capturing:
            # CUDA-graph capture (eags_raster.graph): no event, no host wait, nothing allocated outside
            # the graph's pool.  Capacities come from the eager warm-up steps; every replay refreshes
            # a pinned status block the owner of the graph checks (K <= capacity, no bucket overflow).
            cap_ctx = _capture_ctx
            if cap_ctx is None:
                raise RuntimeError("eags_raster: capture the rasterizer through eags_raster.graph.GraphedIteration "
                                   "(it warms up the capacities and owns the pinned status blocks)")
            key = (idx, W, H)
            cap = _capacity.get(idx, 0)
            if cap <= 0:
                raise RuntimeError("eags_raster: run at least one eager step before capturing a CUDA graph")
            G = ((W + 15) // 16) * ((H + 15) // 16)
            tile_cap = _tile_cap.get(key, 0) if _bucketed else 0
            if tile_cap * G > _MAX_BUCKET_SLOTS:
                tile_cap = 0
            status, sticky = cap_ctx.take_status(cap, device)
            binning = torch.empty(_sizes(lib, "eags_binning_state_bytes_bucketed", cap, int(with_keys), W, H, tile_cap), **u8)
            _lib.check(lib.eags_forward_geom_bucketed(
                C.byref(view), C.byref(g), geom.data_ptr(), image.data_ptr(), radii.data_ptr(),
                status.data_ptr(), binning.data_ptr(), cap, int(with_keys), tile_cap, sticky.data_ptr(), stream),
                "eags_forward_geom_bucketed")
            _lib.check(lib.eags_forward_render_bucketed(
                C.byref(view), C.byref(g), radii.data_ptr(), geom.data_ptr(), image.data_ptr(),
                binning.data_ptr(), cap, int(with_keys), tile_cap, C.byref(imgs), stream),
                "eags_forward_render_bucketed")
            del keep
            state = RasterState(P, M, W, H, variant, cap, bool(with_keys), -1, radii, geom, image, binning)
            return color, depth, third, radii, n_touched, state
        if mode == "optimistic" and P > 0:
            # one pass: stage 1 (+ key scatter into per-tile buckets once a bound is known) and stage 2 are
            # enqueued back to back; the host waits for the stage-1 event only
            key = (idx, W, H)
            G = ((W + 15) // 16) * ((H + 15) // 16)
            tile_cap = _tile_cap.get(key, 0) if _bucketed else 0
            if tile_cap * G > _MAX_BUCKET_SLOTS:
                tile_cap = 0
            cap = _capacity.get(idx, 0) or _round_capacity(max(4 * P, _MIN_CAPACITY))
            status = _pinned_status()
            binning = torch.empty(_sizes(lib, "eags_binning_state_bytes_bucketed", cap, int(with_keys), W, H, tile_cap), **u8)
            _lib.check(lib.eags_forward_geom_bucketed(
                C.byref(view), C.byref(g), geom.data_ptr(), image.data_ptr(), radii.data_ptr(),
                status.data_ptr(), binning.data_ptr(), cap, int(with_keys), tile_cap, None, stream),
                "eags_forward_geom_bucketed")
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(device))
            _lib.check(lib.eags_forward_render_bucketed(
                C.byref(view), C.byref(g), radii.data_ptr(), geom.data_ptr(), image.data_ptr(),
                binning.data_ptr(), cap, int(with_keys), tile_cap, C.byref(imgs), stream),
                "eags_forward_render_bucketed")
            ev.synchronize()
            K = int(status[0])
            s32 = status.view(torch.int32)
            max_tile, overflow = int(s32[7]), int(s32[8])
            if K > cap or overflow:
                # nothing was rendered (every stage-2 kernel exited): re-issue stage 2 on the exact path
                cap = max(cap, _round_capacity(int(K * 1.25)))
                binning = torch.empty(_sizes(lib, "eags_binning_state_bytes", cap, int(with_keys)), **u8)
                _lib.check(lib.eags_forward_render(C.byref(view), C.byref(g), radii.data_ptr(),
                                                   geom.data_ptr(), image.data_ptr(), binning.data_ptr(), cap,
                                                   int(with_keys), C.byref(imgs), stream), "eags_forward_render")
            _capacity[idx] = max(_capacity.get(idx, 0), _round_capacity(int(K * 1.25)))
            _tile_cap[key] = max(_tile_cap.get(key, 0), (int(max_tile * 1.25) + 64 + 63) & ~63)
            del keep
            state = RasterState(P, M, W, H, variant, cap, bool(with_keys), K, radii, geom, image, binning)
            return color, depth, third, radii, n_touched, state
        if mode == "nosync":
            _check_pending(idx)
            slot = _nosync_slot()
        else:
            slot = _pinned_slot()
        _lib.check(lib.eags_forward_geom(C.byref(view), C.byref(g), geom.data_ptr(), image.data_ptr(),
                                         radii.data_ptr(), slot.data_ptr(), stream), "eags_forward_geom")
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(device))

        def render(cap):
            b = torch.empty(_sizes(lib, "eags_binning_state_bytes", cap, int(with_keys)), **u8)
            _lib.check(lib.eags_forward_render(C.byref(view), C.byref(g), radii.data_ptr(),
                                               geom.data_ptr(), image.data_ptr(), b.data_ptr(), cap,
                                               int(with_keys), C.byref(imgs), stream),
                       "eags_forward_render")
            return b

        if mode == "exact":
            ev.synchronize()
            K = int(slot[0])
            cap = max(K, 1)
            binning = render(cap)
        else:
            cap = _capacity.get(idx, 0) or _round_capacity(max(4 * P, _MIN_CAPACITY))
            binning = render(cap)
            if mode == "optimistic":
                ev.synchronize()          # waits for stage 1 only; the render keeps the GPU busy
                K = int(slot[0])
                if K > cap:               # render kernels exited untouched: re-issue with room
                    cap = _round_capacity(int(K * 1.25))
                    binning = render(cap)
                _capacity[idx] = max(_capacity.get(idx, 0), _round_capacity(int(K * 1.25)))
            else:
                K = -1
                pend = getattr(_tls, "pending", None) or []
                pend.append((idx, ev, slot, cap))
                _tls.pending = pend
                _capacity[idx] = cap
        del keep
        state = RasterState(P, M, W, H, variant, cap, bool(with_keys), K, radii, geom, image, binning)
    return color, depth, third, radii, n_touched, state


def rasterize_backward(settings, state: RasterState, means3D, opacities, shs, colors_precomp, scales,
                       rotations, cov3D_precomp, dL_dcolor, dL_ddepth, dL_dthird=None,
                       want_shs=True, want_tau=False, pose_only=False):
    """Returns dict(means3D, means2D, opacities, shs | colors_precomp, scales, rotations |
    cov3D_precomp, tau).  ``pose_only``: only ``tau`` is produced -- no per-Gaussian gradient array is
    allocated or written (the tracking fast path, eags_raster.tracking)."""
    lib = _lib.load()
    device = state.radii.device
    P, M, H, W = state.P, state.M, state.H, state.W
    with torch.cuda.device(device):
        f32 = dict(dtype=torch.float32, device=device)
        means3D = _prep(means3D, device)
        opacities = _prep(opacities, device)
        shs = _prep(shs, device)
        colors_precomp = _prep(colors_precomp, device)
        scales = _prep(scales, device)
        rotations = _prep(rotations, device)
        cov3D_precomp = _prep(cov3D_precomp, device)
        dL_dcolor = _prep(dL_dcolor, device)
        dL_ddepth = _prep(dL_ddepth, device)
        dL_dthird = _prep(dL_dthird, device)
        if dL_dcolor is None:
            dL_dcolor = torch.zeros(3, H, W, **f32)
        if dL_ddepth is None:
            dL_ddepth = torch.zeros(1, H, W, **f32)
        arena = _arena.get(device.index) or {}
        route = _peer_route.get(device.index) if not pose_only else None
        route_final = True
        if route is not None:
            ex = route.ex
            if ex.P != P or ex.device != device or scales is None or rotations is None:
                raise ValueError("PeerGradExchange.route(): the exchange was built for another submap "
                                 f"(P {ex.P} on {ex.device}) or the inputs use cov3D_precomp")
            route_final = route.begin_pass()

        def buf(name, *shape):
            t = arena.get(name)
            if t is not None:
                if t.dtype != torch.float32 or not t.is_contiguous() or t.device != device or \
                        t.numel() != torch.Size(shape).numel():
                    raise ValueError(f"grad_arena['{name}'] must be a contiguous fp32 tensor of shape {shape} on {device}")
                return t.view(*shape)
            return torch.empty(*shape, **f32)

        routed = route is not None     # the four routed rows are written through the route, not through `out`
        out = dict(
            means3D=None if routed else buf("means3D", P, 3),
            means2D=buf("means2D", P, 3),
            opacities=None if routed else buf("opacities", P, 1),
            shs=buf("shs", P, M, 3) if (shs is not None and want_shs) else None,
            colors_precomp=buf("colors_precomp", P, 3) if colors_precomp is not None else None,
            scales=None if routed else (buf("scales", P, 3) if scales is not None else None),
            rotations=None if routed else (buf("rotations", P, 4) if rotations is not None else None),
            cov3D_precomp=buf("cov3D_precomp", P, 6) if cov3D_precomp is not None else None,
            tau=torch.empty(6, **f32) if want_tau else None,
        ) if not pose_only else dict(
            means3D=None, means2D=None, opacities=None, shs=None, colors_precomp=None, scales=None,
            rotations=None, cov3D_precomp=None, tau=torch.empty(6, **f32))
        stream = torch.cuda.current_stream(device).cuda_stream
        view, keep = _make_view(settings, state.variant, device)
        g = _make_gaussians(P, M, means3D, opacities, shs, colors_precomp, scales, rotations,
                            cov3D_precomp)
        gin = EagsImageGrads(dL_dcolor.data_ptr(), dL_ddepth.data_ptr(), _ptr(dL_dthird))
        gout = EagsInputGrads(_ptr(out["means3D"]), _ptr(out["means2D"]), _ptr(out["opacities"]),
                              _ptr(out["shs"]), _ptr(out["colors_precomp"]), _ptr(out["scales"]),
                              _ptr(out["rotations"]), _ptr(out["cov3D_precomp"]), _ptr(out["tau"]))
        scratch = torch.empty(_sizes(lib, "eags_backward_scratch_bytes", P), dtype=torch.uint8, device=device)
        if route is not None and route_final:
            # sharded mapping: xyz / scale / rotation / opacity gradients (plus the local sums of the
            # rank's earlier views of this step) are scattered to their owner ranks by preprocess_bwd
            # and come back reduced over all ranks (eags_raster.peer)
            _lib.check(lib.eags_backward_routed(C.byref(view), C.byref(g), state.radii.data_ptr(),
                                                state.geom.data_ptr(), state.image.data_ptr(),
                                                state.binning.data_ptr(), state.capacity, C.byref(gin),
                                                C.byref(gout), scratch.data_ptr(),
                                                C.byref(ex.route_struct(route.rank, route.pass_accum())), stream),
                       "eags_backward_routed")
            if route.optimizer is not None:
                # fused owner-side step: sum, activation backward, regulariser, Adam, parameter all-gather
                # (eags_raster.peer.PeerAdam); the raw parameters were updated in place on every replica
                if route.do_reduce:
                    route.optimizer.step(route.scale, route.rank)
                out.update(means3D=None, scales=None, rotations=None, opacities=None)
            else:
                if route.do_reduce:
                    ex.reduce(route.scale, route.rank)
                out.update(ex.reduced_views(route.rank))
        elif route is not None:
            # an earlier view of this rank's share of the step: summed into the exchange's local buffer by the
            # same routed preprocess_bwd (a one-rank route onto that buffer).  The four routed inputs get no
            # gradient from this pass: the rank's last pass of the step returns the sum over all views of all
            # ranks (and the chain rule above them runs once)
            _lib.check(lib.eags_backward_routed(C.byref(view), C.byref(g), state.radii.data_ptr(),
                                                state.geom.data_ptr(), state.image.data_ptr(),
                                                state.binning.data_ptr(), state.capacity, C.byref(gin),
                                                C.byref(gout), scratch.data_ptr(), C.byref(route.pass_route()),
                                                stream), "eags_backward_routed")
        else:
            _lib.check(lib.eags_backward(C.byref(view), C.byref(g), state.radii.data_ptr(),
                                         state.geom.data_ptr(), state.image.data_ptr(),
                                         state.binning.data_ptr(), state.capacity, C.byref(gin),
                                         C.byref(gout), scratch.data_ptr(), stream), "eags_backward")
        if _kernel_stats:
            off = int(lib.eags_backward_stats_offset(P))
            st64 = scratch[off:off + 16].view(torch.int64)
            out["stats"] = {"warp_iters": int(st64[0]), "lane_pairs": int(st64[1])}
        del keep
    return out


def state_array(state: RasterState, name: str) -> torch.Tensor:
    """Copy one named array out of the opaque state buffers (parity tests / debugging)."""
    if name in ("xy", "rgbd", "conic_opacity"):
        # views derived from the per-Gaussian splat records (the conic is stored pre-scaled for the blend
        # kernels: undo the -log2(e)/2, -log2(e), -log2(e)/2 factors)
        rec = state_array(state, "splat_rec")
        if name == "xy":
            return rec[:, 0:2].contiguous()
        if name == "rgbd":
            return torch.cat([rec[:, 8:11], rec[:, 3:4]], dim=1).contiguous()
        L = 1.4426950408889634
        return torch.stack([rec[:, 4] / (-0.5 * L), rec[:, 5] / (-L), rec[:, 6] / (-0.5 * L), rec[:, 7]], dim=1)
    lib = _lib.load()
    ptr, cnt, eb = C.c_void_p(0), C.c_int64(0), C.c_int32(0)
    _lib.check(lib.eags_state_view(name.encode(), state.P, state.W, state.H, state.capacity,
                                   int(state.with_keys), state.geom.data_ptr(), state.image.data_ptr(),
                                   state.binning.data_ptr(), C.byref(ptr), C.byref(cnt), C.byref(eb)),
               "eags_state_view")
    for buf in (state.geom, state.image, state.binning):
        off = (ptr.value or 0) - buf.data_ptr()
        if 0 <= off < buf.numel():
            raw = buf[off: off + cnt.value * eb.value].clone()
            break
    else:
        raise RuntimeError("state array outside the state buffers")
    dt = {"splat_rec": (torch.float32, (-1, 12)), "tiles_touched": (torch.int32, (-1,)),
          "clamped": (torch.uint8, (-1,)), "tile_count": (torch.int32, (-1,)),
          "ranges": (torch.int32, (-1, 2)), "tile_max_contrib": (torch.int32, (-1,)),
          "final_T": (torch.float32, (state.H, state.W)), "n_contrib": (torch.int32, (state.H, state.W)),
          "point_list": (torch.int32, (-1,)), "point_list_keys": (torch.int64, (-1,)),
          "records": (torch.float32, (-1, 12)), "ctrl": (torch.int64, (-1,))}[name]
    return raw.view(dt[0]).reshape(dt[1])


# ---- autograd bridge -------------------------------------------------------------------------------
class _Rasterize(torch.autograd.Function):
    """Autograd node shared by both API flavours (upstream: ``_RasterizeGaussians``)."""

    @staticmethod
    def forward(ctx, means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                cov3D_precomp, theta, rho, settings, variant):
        color, depth, third, radii, n_touched, state = rasterize_forward(
            settings, variant, means3D, opacities, shs, colors_precomp, scales, rotations,
            cov3D_precomp, want_n_touched=(variant == _lib.VARIANT_B))
        ctx.settings = settings
        ctx.state = state
        ctx.has = [t is not None and t.numel() > 0 for t in
                   (shs, colors_precomp, scales, rotations, cov3D_precomp)]
        ctx.want_theta, ctx.want_rho = theta is not None, rho is not None
        ctx.want_tau = ctx.want_theta or ctx.want_rho
        ctx.save_for_backward(*[t if t is not None else torch.empty(0) for t in
                                (means3D, opacities, shs, colors_precomp, scales, rotations,
                                 cov3D_precomp)])
        ctx.mark_non_differentiable(radii)
        if n_touched is not None:
            ctx.mark_non_differentiable(n_touched)
            return color, depth, third, radii, n_touched
        return color, depth, third, radii

    @staticmethod
    def backward(ctx, g_color, g_depth, g_third, *unused):
        means3D, opacities, shs, colors_precomp, scales, rotations, cov3D = [
            t if t.numel() > 0 else None for t in ctx.saved_tensors]
        st = ctx.state
        needs = ctx.needs_input_grad
        if st.P == 0:
            z = lambda t: None if t is None else torch.zeros_like(t)
            z3 = lambda want: torch.zeros(3, device=st.radii.device) if want else None
            return (z(means3D), None, z(shs), z(colors_precomp), z(opacities), z(scales), z(rotations),
                    z(cov3D), z3(ctx.want_theta), z3(ctx.want_rho), None, None)
        # variant B upstream drops the gradient of the opacity image; variant A carries dL/dalpha
        g_third = g_third if st.variant == _lib.VARIANT_A else None
        # registration (src/gsr/solver.py) optimises the pose only: no per-Gaussian gradient is written then
        pose_only = ctx.want_tau and not any(needs[:8])
        out = rasterize_backward(ctx.settings, st, means3D, opacities, shs, colors_precomp, scales,
                                 rotations, cov3D, g_color, g_depth, g_third,
                                 want_shs=bool(needs[2]), want_tau=ctx.want_tau, pose_only=pose_only)
        # a gradient only for the pose slots that were given a tensor (theta alone or rho alone is legal)
        tau = out["tau"]
        g_rho = tau[:3] if (tau is not None and ctx.want_rho) else None
        g_theta = tau[3:] if (tau is not None and ctx.want_theta) else None
        if pose_only:
            return (None, None, None, None, None, None, None, None, g_theta, g_rho, None, None)
        return (out["means3D"], out["means2D"], out["shs"], out["colors_precomp"],
                out["opacities"].view(opacities.shape) if out["opacities"] is not None else None,
                out["scales"], out["rotations"], out["cov3D_precomp"], g_theta, g_rho, None, None)


def rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                        cov3D_precomp, settings, variant, theta=None, rho=None):
    return _Rasterize.apply(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                            cov3D_precomp, theta, rho, settings, variant)
