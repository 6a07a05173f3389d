"""Gradient exchange of sharded mapping over NVLink peer memory (csrc/collective.cu).

:class:`PeerGradExchange` owns the symmetric (peer-mapped) buffers of one replicated submap --
PyTorch's symmetric-memory allocator provides the memory and the handle exchange, nothing else -- and,
while ``route()`` is active, makes the rasterizer's backward pass

1. store each Gaussian's 11 parameter-gradient floats straight into the owner rank's receive slot
   (reduce-scatter fused into ``preprocess_bwd``), and
2. launch ONE ``peer_reduce`` kernel (barrier, sum in rank order, ``x scale``, push to every rank,
   barrier),

so that ``loss.backward()`` returns with the *averaged* gradients of all ranks in ``p.grad`` -- views
of the local ``reduced`` buffer, bit-identical on every rank.  The NCCL path
(:class:`eags_raster.sharded.FusedGradBucket`) stays as the general fallback (several views per rank
per step, or no peer access).

``emulate_world=N`` builds N ranks' buffers on ONE device with the cross-rank waits disabled: the
single-GPU parity test of the exchange (ranks run one after the other).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib

MAX_PEERS = 8


class EagsPeerRoute(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("P", C.c_int64),
                ("recv", C.c_void_p * MAX_PEERS), ("reduced", C.c_void_p * MAX_PEERS),
                ("flags", C.c_void_p * MAX_PEERS), ("accum", C.c_void_p * 4)]


class EagsPeerAdam(C.Structure):
    _fields_ = [("params", C.c_void_p * MAX_PEERS), ("m", C.c_void_p), ("v", C.c_void_p), ("step", C.c_void_p),
                ("lr", C.c_float * 4), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
                ("isotropic_weight", C.c_float)]


class PeerExchangeError(RuntimeError):
    """A bounded cross-rank wait expired: a peer never delivered its gradients for some step."""


def stride_of(P: int) -> int:
    return (int(P) + 3) & ~3


def owner_chunk(P: int, world: int) -> int:
    """Gaussians per owner rank (the last rank may own fewer): ceil(P / world) rounded up to 4."""
    return (((int(P) + world - 1) // world) + 3) & ~3


class PeerGradExchange:
    def __init__(self, P: int, device, group=None, emulate_world: Optional[int] = None):
        self.P = int(P)
        self.group = group
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("PeerGradExchange needs a CUDA device (no CPU fallback)")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.stride = stride_of(P)
        self.flat = 11 * self.stride
        self.epoch = 0
        lib = _lib.load()
        self.flags_bytes = int(lib.eags_peer_flags_bytes())
        self.emulated = emulate_world is not None
        if self.emulated:
            self.world, self.rank = int(emulate_world), 0
        else:
            import torch.distributed as dist
            self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if not 1 <= self.world <= MAX_PEERS:
            raise ValueError(f"PeerGradExchange supports 1..{MAX_PEERS} ranks of one node, got {self.world}")
        recv_b, red_b = self.world * self.flat * 4, self.flat * 4
        self._off_reduced = (recv_b + 255) & ~255
        self._off_flags = (self._off_reduced + red_b + 255) & ~255
        # raw parameter replica (PeerAdam: owners store updated parameters straight into every replica)
        self._off_params = (self._off_flags + self.flags_bytes + 255) & ~255
        total = self._off_params + red_b
        if self.emulated:
            self._bufs = [torch.zeros(total, dtype=torch.uint8, device=self.device) for _ in range(self.world)]
            bases = [b.data_ptr() for b in self._bufs]
            self._handle = None
        else:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem
            pg = group if group is not None else dist.group.WORLD
            with torch.cuda.device(self.device):
                buf = symm_mem.empty(total, dtype=torch.uint8, device=self.device)
                self._handle = symm_mem.rendezvous(buf, pg)
            buf.zero_()
            torch.cuda.synchronize(self.device)
            dist.barrier(group)          # every rank's flags are zero before anyone signals
            self._bufs = [buf]
            bases = [int(p) for p in self._handle.buffer_ptrs]
            if len(bases) != self.world or bases[self.rank] != buf.data_ptr():
                raise RuntimeError("symmetric-memory rendezvous returned unexpected peer pointers")
        self._bases = bases
        self._routes = [self._make_route(r) for r in (range(self.world) if self.emulated else [self.rank])]
        self._sums = None

    def _make_route(self, rank: int) -> EagsPeerRoute:
        r = EagsPeerRoute()
        r.rank, r.world, r.P = rank, self.world, self.P
        for p in range(self.world):
            r.recv[p] = self._bases[p]
            r.reduced[p] = self._bases[p] + self._off_reduced
            r.flags[p] = self._bases[p] + self._off_flags
        return r

    # ---- views of the local result --------------------------------------------------------------
    def _local_buf(self, rank: Optional[int] = None) -> torch.Tensor:
        return self._bufs[rank if self.emulated else 0]

    def reduced_views(self, rank: Optional[int] = None):
        """dict(means3D (P,3), scales (P,3), rotations (P,4), opacities (P,1)): views of the result."""
        rank = self.rank if rank is None else rank
        flat = self._local_buf(rank)[self._off_reduced:self._off_reduced + self.flat * 4].view(torch.float32)
        s, P = self.stride, self.P
        return {"means3D": flat[0:3 * P].view(P, 3), "scales": flat[3 * s:3 * s + 3 * P].view(P, 3),
                "rotations": flat[6 * s:6 * s + 4 * P].view(P, 4), "opacities": flat[10 * s:10 * s + P].view(P, 1)}

    def recv_views(self, owner: int, src: int):
        """dict(means3D, scales, rotations, opacities): views of slot ``src`` inside rank ``owner``'s receive area
        (emulated worlds only -- tests write known gradients here)."""
        if not self.emulated:
            raise RuntimeError("recv_views is for emulate_world exchanges")
        flat = self._bufs[owner][src * self.flat * 4:(src + 1) * self.flat * 4].view(torch.float32)
        return self._carve(flat)

    def error_flag(self, rank: Optional[int] = None) -> int:
        """Non-zero (the first epoch that failed... sticky maximum) if a bounded cross-rank wait expired on
        this rank.  Synchronises the host with the device (one 4-byte read)."""
        rank = self.rank if rank is None else rank
        words = self._local_buf(rank)[self._off_flags:self._off_flags + self.flags_bytes].view(torch.int32)
        return int(words[2 * MAX_PEERS + 1])

    def check(self, rank: Optional[int] = None) -> None:
        """Raise :class:`PeerExchangeError` if any exchange since the last check lost a peer.  The step that
        failed delivered NaN gradients (``reduce``) or skipped its update (``PeerAdam``) -- never a partial
        sum -- so the caller can stop, fall back to NCCL, or re-run the step."""
        e = self.error_flag(rank)
        if e:
            raise PeerExchangeError(f"peer gradient exchange: rank {self.rank if rank is None else rank} waited "
                                    f"30 s for a peer at exchange #{e}; that step's result was discarded")

    def param_views(self, rank: Optional[int] = None):
        """dict(xyz (P,3), scaling (P,3), rotation (P,4), opacity (P,1)): views of this rank's RAW parameter
        replica inside the peer-mapped buffer (the segment layout of the gradient buffers)."""
        rank = self.rank if rank is None else rank
        flat = self._local_buf(rank)[self._off_params:self._off_params + self.flat * 4].view(torch.float32)
        s, P = self.stride, self.P
        return {"xyz": flat[0:3 * P].view(P, 3), "scaling": flat[3 * s:3 * s + 3 * P].view(P, 3),
                "rotation": flat[6 * s:6 * s + 4 * P].view(P, 4), "opacity": flat[10 * s:10 * s + P].view(P, 1)}

    # ---- the routed backward ----------------------------------------------------------------------
    def route(self, scale: float = 1.0, rank: Optional[int] = None, reduce: bool = True, accum=None,
              local_views: int = 1, optimizer: "Optional[PeerAdam]" = None):
        """Context manager: the next rasterizer backward scatters to the owners and (``reduce``) runs the
        peer reduction with ``scale`` (1 / number of views).  ``rank`` selects the emulated rank.

        ``accum`` -- dict(means3D (P,3), scales (P,3), rotations (P,4), opacities (P,1)) of contiguous
        fp32 tensors on this device: the rank's gradient sums over the views it ALREADY rendered this
        step (plain backward passes).  They are added to the routed rows in ``preprocess_bwd``, so with
        several views per rank per step only the last backward goes over NVLink.

        ``local_views=k`` does that bookkeeping here: the context then expects exactly k rasterizer
        backward passes; the first k-1 run the plain backward, are summed into a buffer of this object and
        hand NO gradient to the four routed inputs; the k-th is routed with that sum added and returns
        the all-rank result, so the activations above the rasterizer back-propagate once per step.

        ``optimizer`` (:class:`PeerAdam`) -- the routed pass ends with the fused owner-side optimiser step
        instead of the gradient reduction: NO gradient is returned for the four routed inputs (the raw
        parameters inside the exchange are updated in place on every replica)."""
        return _RouteCtx(self, float(scale), self.rank if rank is None else rank, reduce, accum, local_views,
                         optimizer)

    def _local_sums(self) -> torch.Tensor:
        """The flat 11 x stride fp32 buffer of ``route(local_views > 1)``: this rank's gradient sums over the
        views it already rendered this step (made on first use)."""
        if self._sums is None:
            self._sums = torch.zeros(self.flat, dtype=torch.float32, device=self.device)
            self._self_routes = None
        return self._sums

    def _self_route(self, first: bool) -> EagsPeerRoute:
        """A one-rank route whose only "owner slot" is the local sums buffer: a non-final pass of
        ``route(local_views > 1)`` goes through the same routed ``preprocess_bwd`` and leaves
        ``sums = new`` (first pass) or ``sums = sums + new`` (accum = the buffer itself: every CTA reads
        and writes only its own rows) -- no scratch copy and no separate add pass over 44 B per Gaussian."""
        acc = self._local_sums()
        if self._self_routes is None:
            routes = []
            for with_acc in (False, True):
                r = EagsPeerRoute()
                r.rank, r.world, r.P = 0, 1, self.P
                r.recv[0] = r.reduced[0] = acc.data_ptr()
                r.flags[0] = self._bases[self.rank if not self.emulated else 0] + self._off_flags   # never touched
                if with_acc:
                    for k, t in enumerate(self._carve(acc).values()):
                        r.accum[k] = t.data_ptr()
                routes.append(r)
            self._self_routes = routes
        return self._self_routes[0 if first else 1]

    def _carve(self, flat: torch.Tensor):
        s, P = self.stride, self.P
        return {"means3D": flat[0:3 * P].view(P, 3), "scales": flat[3 * s:3 * s + 3 * P].view(P, 3),
                "rotations": flat[6 * s:6 * s + 4 * P].view(P, 4), "opacities": flat[10 * s:10 * s + P].view(P, 1)}

    def route_struct(self, rank: Optional[int] = None, accum=None) -> EagsPeerRoute:
        rank = self.rank if rank is None else rank
        base = self._routes[rank if self.emulated else 0]
        if accum is None:
            return base
        r = EagsPeerRoute()
        C.memmove(C.byref(r), C.byref(base), C.sizeof(EagsPeerRoute))
        shapes = {"means3D": (self.P, 3), "scales": (self.P, 3), "rotations": (self.P, 4), "opacities": (self.P, 1)}
        for k, (name, shape) in enumerate(shapes.items()):
            t = accum.get(name)
            if t is None or tuple(t.shape) != shape or t.dtype != torch.float32 or not t.is_contiguous() \
                    or t.device != self.device:
                raise ValueError(f"route(accum=...): '{name}' must be a contiguous fp32 {shape} tensor on {self.device}")
            r.accum[k] = t.data_ptr()
        return r

    def reduce(self, scale: float = 1.0, rank: Optional[int] = None, epoch: Optional[int] = None) -> None:
        """Launch the peer reduction on the current stream (all ranks must call it once per step)."""
        lib = _lib.load()
        if epoch is None:
            epoch = 0     # device-side epoch counter (also valid inside a CUDA-graph replay)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(lib.eags_peer_reduce(C.byref(self.route_struct(rank)), epoch & 0xffffffff, float(scale),
                                        0 if self.emulated else 1, stream), "eags_peer_reduce")


class PeerAdam:
    """The mapping optimiser of sharded mapping, fused into the gradient exchange (csrc/collective.cu
    ``peer_adam_kernel``; include/eags_raster.h ``eags_peer_adam_step``).

    Holds the RAW parameters of the reference's ``GaussianModel`` (``_xyz``, ``_scaling`` = log scale,
    ``_rotation`` = un-normalised quaternion, ``_opacity`` = logit; gaussian_model.py:60-81) inside the
    exchange's peer-mapped buffer and the Adam moments of the Gaussians THIS rank owns (1 / world of the
    optimiser state).  ``activated()`` gives the tensors the rasterizer takes; after the step's last
    backward pass (``exchange.route(..., optimizer=self)``) the owner of each Gaussian sums the ranks'
    gradients, back-propagates through exp / normalize / sigmoid, adds the isotropic regulariser's gradient
    (mapper.py:166), takes the Adam step of gaussian_model.py:341-356 and stores the new parameters into every
    replica.  At world size 1 (or ``emulate_world=1``) this is the reference's single-GPU optimiser step.
    """

    GROUPS = ("xyz", "scaling", "rotation", "opacity")

    def __init__(self, exchange: PeerGradExchange, lr=None, betas=(0.9, 0.999), eps: float = 1e-15,
                 isotropic_weight: float = 1.0):
        # learning rates of src/entities/arguments.py:57-64
        self.lr = dict(xyz=1e-4, scaling=5e-3, rotation=1e-3, opacity=5e-2)
        self.lr.update(lr or {})
        self.ex, self.betas, self.eps, self.isotropic_weight = exchange, betas, float(eps), float(isotropic_weight)
        ranks = range(exchange.world) if exchange.emulated else [exchange.rank]
        chunk = owner_chunk(exchange.P, exchange.world)
        dev = exchange.device
        self._state = {r: (torch.zeros(11 * chunk, dtype=torch.float32, device=dev),
                           torch.zeros(11 * chunk, dtype=torch.float32, device=dev),
                           torch.zeros(1, dtype=torch.int32, device=dev)) for r in ranks}
        self._structs = {r: self._make_struct(r) for r in ranks}

    def _make_struct(self, rank: int) -> EagsPeerAdam:
        ex, o = self.ex, EagsPeerAdam()
        for p in range(ex.world):
            o.params[p] = ex._bases[p] + ex._off_params
        m, v, step = self._state[rank]
        o.m, o.v, o.step = m.data_ptr(), v.data_ptr(), step.data_ptr()
        for k, name in enumerate(self.GROUPS):
            o.lr[k] = float(self.lr[name])
        o.beta1, o.beta2, o.eps, o.isotropic_weight = float(self.betas[0]), float(self.betas[1]), self.eps, self.isotropic_weight
        return o

    def set_lr(self, **lr) -> None:
        """e.g. the xyz schedule of gaussian_model.py:326-333."""
        self.lr.update(lr)
        for r in self._structs:
            self._structs[r] = self._make_struct(r)

    def params(self, rank: Optional[int] = None):
        return self.ex.param_views(rank)

    def load(self, xyz, scaling, rotation, opacity, rank: Optional[int] = None) -> None:
        """Copy initial raw parameters into the replica (every rank loads the same values) and reset the
        optimiser state."""
        pv = self.params(rank)
        with torch.no_grad():
            pv["xyz"].copy_(xyz); pv["scaling"].copy_(scaling)
            pv["rotation"].copy_(rotation); pv["opacity"].copy_(opacity.reshape(-1, 1))
        ranks = [self.ex.rank if rank is None else rank]
        for r in ranks:
            for t in self._state[r]:
                t.zero_()

    def activated(self, rank: Optional[int] = None):
        """dict(means3D, scales, rotations, opacities) for the rasterizer (gaussian_model.py:103-122);
        no autograd graph: the chain rule through the activations is part of the fused step."""
        pv = self.params(rank)
        lib, dev = _lib.load(), self.ex.device
        P = int(pv["xyz"].shape[0])
        with torch.no_grad(), torch.cuda.device(dev):
            f32 = dict(dtype=torch.float32, device=dev)
            s, q, o = torch.empty(P, 3, **f32), torch.empty(P, 4, **f32), torch.empty(P, 1, **f32)
            # exp / F.normalize / sigmoid in one kernel (eags_gaussian_activations)
            _lib.check(lib.eags_gaussian_activations(P, pv["scaling"].data_ptr(), pv["rotation"].data_ptr(),
                                                     pv["opacity"].data_ptr(), s.data_ptr(), q.data_ptr(),
                                                     o.data_ptr(), torch.cuda.current_stream(dev).cuda_stream),
                       "eags_gaussian_activations")
        return {"means3D": pv["xyz"], "scales": s, "rotations": q, "opacities": o}

    # ---- the optimiser surgery of a growing / shrinking submap (gaussian_model.py:501-584) ----------------
    def _full_moments(self):
        """(m, v): the (11, P) first / second moments of ALL Gaussians (rows: xyz 0-2, scaling 3-5, rotation 6-9,
        opacity 10), assembled from the owners' slices; identical on every rank."""
        ex = self.ex
        chunk = owner_chunk(ex.P, ex.world)
        if ex.emulated:
            m = torch.cat([self._state[r][0].view(11, chunk) for r in range(ex.world)], dim=1)
            v = torch.cat([self._state[r][1].view(11, chunk) for r in range(ex.world)], dim=1)
        elif ex.world == 1:
            m, v = self._state[ex.rank][0].view(11, chunk), self._state[ex.rank][1].view(11, chunk)
        else:
            import torch.distributed as dist
            m_loc, v_loc, _ = self._state[ex.rank]
            out = torch.empty(ex.world, 2, 11 * chunk, dtype=torch.float32, device=ex.device)
            dist.all_gather_into_tensor(out, torch.stack([m_loc, v_loc]), group=ex.group)
            m = out[:, 0].reshape(ex.world, 11, chunk).permute(1, 0, 2).reshape(11, ex.world * chunk)
            v = out[:, 1].reshape(ex.world, 11, chunk).permute(1, 0, 2).reshape(11, ex.world * chunk)
        return m[:, :ex.P], v[:, :ex.P]

    def resized(self, keep: Optional[torch.Tensor] = None, new: Optional[dict] = None) -> "PeerAdam":
        """A new PeerAdam (on a new exchange of the new size) after the reference's two kinds of optimiser
        surgery, applied in this order:

        ``keep`` -- bool (P,) mask of the Gaussians that stay: ``prune_points`` / ``_prune_optimizer``
        (gaussian_model.py:501-533; mapper.py:206-207, 222-223): parameters AND Adam moments of the kept
        Gaussians are carried over, the step count is kept;
        ``new`` -- dict(xyz (n,3), scaling (n,3), rotation (n,4), opacity (n,1)) of RAW parameters appended with
        zero moments: ``densification_postfix`` / ``cat_tensors_to_optimizer`` (gaussian_model.py:535-583).

        Both arguments must be identical on every rank (they are functions of the replicated parameters); every
        rank calls this collectively (one all-gather of the moments, one symmetric-memory rendezvous).  The
        ownership ranges move with P, so the moments are re-sliced; learning rates, betas, eps carry over.
        Not a per-step operation: the reference prunes three times per ``optimize_submap`` and grows once per
        mapped frame."""
        ex = self.ex
        with torch.no_grad():
            m, v = self._full_moments()
            pv = ex.param_views(0 if ex.emulated else None)
            par = {k: pv[k] for k in self.GROUPS}
            if keep is not None:
                keep = keep.reshape(-1).to(device=ex.device, dtype=torch.bool)
                if keep.numel() != ex.P:
                    raise ValueError(f"resized(keep=...): mask of {keep.numel()} entries for {ex.P} Gaussians")
                m, v = m[:, keep], v[:, keep]
                par = {k: t[keep] for k, t in par.items()}
            if new is not None:
                n = int(new["xyz"].shape[0])
                add = {k: new[k].to(device=ex.device, dtype=torch.float32).reshape(n, -1) for k in self.GROUPS}
                par = {k: torch.cat([par[k], add[k]]) for k in self.GROUPS}
                z = torch.zeros(11, n, dtype=torch.float32, device=ex.device)
                m, v = torch.cat([m, z], dim=1), torch.cat([v, z], dim=1)
            P2 = int(par["xyz"].shape[0])
            if P2 == 0:
                raise ValueError("resized(): no Gaussian left")
            steps = {r: self._state[r][2].clone() for r in self._state}
            ex2 = PeerGradExchange(P2, ex.device, emulate_world=ex.world) if ex.emulated \
                else PeerGradExchange(P2, ex.device, group=ex.group)
            out = PeerAdam(ex2, lr=dict(self.lr), betas=self.betas, eps=self.eps, isotropic_weight=self.isotropic_weight)
            chunk2 = owner_chunk(P2, ex2.world)
            for r in out._state:
                out.load(par["xyz"], par["scaling"], par["rotation"], par["opacity"], rank=r if ex2.emulated else None)
                g0, g1 = min(P2, r * chunk2), min(P2, (r + 1) * chunk2)
                out._state[r][0].view(11, chunk2)[:, :g1 - g0].copy_(m[:, g0:g1])
                out._state[r][1].view(11, chunk2)[:, :g1 - g0].copy_(v[:, g0:g1])
                out._state[r][2].copy_(steps[r])
        return out

    def state_dict(self) -> dict:
        """A COPY of this rank's optimiser state (moments of the Gaussians it owns + step count): the
        checkpoint of mapper.py:189-195 (``optimizer.state_dict()``), restorable with ``load_state_dict``."""
        return {"P": self.ex.P, "world": self.ex.world,
                "state": {r: tuple(t.clone() for t in st) for r, st in self._state.items()}}

    def load_state_dict(self, sd: dict) -> None:
        if sd["P"] != self.ex.P or sd["world"] != self.ex.world:
            raise ValueError("load_state_dict: the checkpoint belongs to a submap of another size (taken before a "
                             "prune / grow)")
        with torch.no_grad():
            for r, st in sd["state"].items():
                for dst, src in zip(self._state[r], st):
                    dst.copy_(src)

    def steps_taken(self, rank: Optional[int] = None) -> int:
        return int(self._state[self.ex.rank if rank is None else rank][2])

    def step(self, scale: float = 1.0, rank: Optional[int] = None) -> None:
        """Launch the fused step on the current stream (normally called by the routed backward)."""
        lib, ex = _lib.load(), self.ex
        rank = ex.rank if rank is None else rank
        stream = torch.cuda.current_stream(ex.device).cuda_stream
        _lib.check(lib.eags_peer_adam_step(C.byref(ex.route_struct(rank)), C.byref(self._structs[rank]), 0,
                                           float(scale), 0 if ex.emulated else 1, stream), "eags_peer_adam_step")


class _RouteCtx:
    def __init__(self, ex: PeerGradExchange, scale: float, rank: int, reduce: bool, accum=None,
                 local_views: int = 1, optimizer: Optional[PeerAdam] = None):
        if local_views < 1:
            raise ValueError("route(local_views=...) must be >= 1")
        if accum is not None and local_views != 1:
            raise ValueError("route(): pass either accum= (caller-owned sums) or local_views= (summed here)")
        if optimizer is not None and optimizer.ex is not ex:
            raise ValueError("route(optimizer=...): the optimiser belongs to another exchange")
        self.ex, self.scale, self.rank, self.do_reduce, self.accum = ex, scale, rank, reduce, accum
        self.local_views, self.passes, self.optimizer = int(local_views), 0, optimizer

    # ---- several views per rank per step (called by rasterize_backward) ---------------------------
    def begin_pass(self) -> bool:
        """True if this backward pass is the rank's last of the step (the routed one)."""
        if self.passes >= self.local_views:
            raise RuntimeError(f"route(local_views={self.local_views}): more rasterizer backward passes than announced")
        self.passes += 1
        return self.passes == self.local_views

    def pass_route(self) -> EagsPeerRoute:
        """The route of a NON-final pass: into the local sums (stored by the first pass, added to after)."""
        return self.ex._self_route(first=self.passes == 1)

    def pass_accum(self):
        if self.accum is not None:
            return self.accum
        return self.ex._carve(self.ex._local_sums()) if self.local_views > 1 else None

    def __enter__(self):
        from . import rasterize
        idx = self.ex.device.index
        self.prev = rasterize._peer_route.get(idx)
        rasterize._peer_route[idx] = self       # per device: another thread's renders elsewhere are unaffected
        return self

    def __exit__(self, *exc):
        from . import rasterize
        idx = self.ex.device.index
        if self.prev is None:
            rasterize._peer_route.pop(idx, None)
        else:
            rasterize._peer_route[idx] = self.prev
        if exc[0] is None and self.passes not in (0, self.local_views):
            raise RuntimeError(f"route(local_views={self.local_views}) saw {self.passes} backward passes: "
                               "the step's gradients were never exchanged")
        return False
