"""Keyframe-batch mapping over several GPUs, and submap placement for parallel loop closure.

The reference optimises a submap on one GPU, one random keyframe view per iteration
(src/entities/mapper.py:143-170), and runs loop-closure registration in one worker thread on one
other device (src/entities/lc.py:60,625).  It has no collective anywhere.
This module is the multi-GPU extension of SURVEY 8(e):

* **sharded mapping** -- the path shards over *views*.  One process per GPU; every rank holds a
  full replica of the submap's parameters and optimiser state, renders its own keyframe view(s) of the
  step, and the parameter gradients are summed with ONE all-reduce over a fused flat buffer
  (:class:`FusedGradBucket`: ``p.grad`` of every parameter is a view into it, so autograd
  accumulates in place and nothing is packed or unpacked), divided by the number of views, and the
  identical optimiser step is taken on every rank.  At world_size 1 with one view per step this is
  exactly the reference's step.
* **parallel loop closure** -- independent submaps are placed round-robin, no collective
  (:func:`assign_submaps`).

Only ``torch.distributed`` is used for the plumbing (NCCL on the GPUs; the host logic is tested with
gloo on CPU, tests/test_sharded_gloo.py).  Nothing here touches the rasterizer kernels; the render
function is whatever the caller passes (``render_gaussian_model`` of the reference, utils.py:120).
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist


def _world() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def _rank() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def shard_views(n_views: int, rank: Optional[int] = None, world: Optional[int] = None) -> List[int]:
    """Indices of the views of one step that ``rank`` renders (round-robin: view v -> rank v % world)."""
    rank = _rank() if rank is None else rank
    world = _world() if world is None else world
    return [v for v in range(int(n_views)) if v % world == rank]


def assign_submaps(n_submaps: int, rank: Optional[int] = None, world: Optional[int] = None) -> List[int]:
    """Independent submap jobs of parallel loop closure handled by ``rank`` (no collective)."""
    return shard_views(n_submaps, rank, world)


class FusedGradBucket:
    """One flat fp32 buffer holding the gradients of all parameters, in the given order.

    ``attach()`` zeroes the buffer (one memset) and points every ``p.grad`` at its slice, so
    ``backward()`` accumulates straight into the bucket; ``all_reduce()`` is then a single
    collective over ``11 * P`` floats for a submap (xyz 3 + scale 3 + rotation 4 + opacity 1).
    """

    def __init__(self, params: Sequence[torch.Tensor]):
        self.params = list(params)
        if not self.params:
            raise ValueError("FusedGradBucket needs at least one parameter")
        dev, dt = self.params[0].device, self.params[0].dtype
        for p in self.params:
            if p.device != dev or p.dtype != dt:
                raise ValueError("all parameters of a bucket must share device and dtype")
        self.numel = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(self.numel, dtype=dt, device=dev)
        self.views, off = [], 0
        for p in self.params:
            self.views.append(self.flat[off:off + p.numel()].view(p.shape))
            off += p.numel()

    def matches(self, params: Sequence[torch.Tensor]) -> bool:
        params = list(params)
        return len(params) == len(self.params) and all(
            a is b and a.shape == v.shape for a, b, v in zip(params, self.params, self.views))

    def attach(self) -> None:
        self.flat.zero_()
        for p, v in zip(self.params, self.views):
            p.grad = v

    def all_reduce(self, divide_by: float = 1.0, async_op: bool = False):
        work = None
        if _world() > 1:
            work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, async_op=async_op)
        if divide_by != 1.0:
            if work is not None and async_op:
                work.wait()
                work = None
            self.flat.mul_(1.0 / float(divide_by))
        return work


def broadcast_parameters(params: Iterable[torch.Tensor], src: int = 0) -> None:
    """Make every replica start from rank ``src``'s values."""
    if _world() == 1:
        return
    with torch.no_grad():
        for p in params:
            dist.broadcast(p.data, src=src)


def consistent_mask(mask: torch.Tensor) -> torch.Tensor:
    """Prune/keep decisions must be identical on all ranks so P stays the same everywhere
    (reference prunes inside the optimisation loop, mapper.py:201-223): a Gaussian is removed only
    if every rank agrees (logical AND via MIN all-reduce)."""
    if _world() == 1:
        return mask
    m = mask.to(torch.int32)
    dist.all_reduce(m, op=dist.ReduceOp.MIN)
    return m.to(torch.bool)


class ShardedMapper:
    """Data-parallel submap optimisation over keyframe views.

    ``loss_fn(view) -> scalar tensor`` renders one view with the *local* replica and returns its
    loss (e.g. the body of the reference's ``optimize_submap`` iteration, mapper.py:152-167).
    ``reg_fn() -> scalar`` is an optional view-independent regulariser (isotropic loss,
    mapper.py:166) added once, on every rank identically, after the all-reduce division.
    """

    def __init__(self, params: Sequence[torch.Tensor], optimizer: Optional[torch.optim.Optimizer],
                 broadcast: bool = True, exchange=None, exchanged: Sequence[torch.Tensor] = (),
                 fused_optimizer=None, check_every: int = 50):
        """``exchange`` (:class:`eags_raster.peer.PeerGradExchange`) moves the gradients of the
        rasterizer's xyz / scale / rotation / opacity inputs over NVLink peer memory inside the backward
        kernel; ``exchanged`` names the parameters that receive gradient ONLY through those four inputs
        (e.g. ``_xyz, _scaling, _rotation, _opacity``, gaussian_model.py:60-79) -- they are left out of the
        NCCL bucket, which then carries just the rest (e.g. the SH features).

        ``fused_optimizer`` (:class:`eags_raster.peer.PeerAdam`) owns xyz / scaling / rotation / opacity
        itself (raw parameters inside the exchange, Adam moments on the owner rank) and takes their step
        inside the exchange kernel; ``params`` / ``optimizer`` then cover only the REST (e.g. SH features;
        both may be empty / None).  ``loss_fn`` must render with ``fused_optimizer.activated()``.

        Every ``check_every`` steps (a host sync) the exchange's error word is read and
        :class:`eags_raster.peer.PeerExchangeError` raised if a peer ever failed to arrive: the failed step
        delivered NaN (reduce) or skipped its update (fused step), never a partial sum."""
        self.optimizer = optimizer
        self.exchange = exchange
        self.fused_optimizer = fused_optimizer
        self.check_every, self._steps = int(check_every), 0
        if fused_optimizer is not None and fused_optimizer.ex is not exchange:
            raise ValueError("ShardedMapper: fused_optimizer must be built on the same exchange=")
        self._exchanged_ids = {id(p) for p in exchanged}
        if exchange is None and self._exchanged_ids:
            raise ValueError("ShardedMapper: exchanged= needs exchange=")
        self.rebuild(params)
        if broadcast:
            broadcast_parameters(list(params))

    def rebuild(self, params: Sequence[torch.Tensor], optimizer: Optional[torch.optim.Optimizer] = None,
                exchange=None, exchanged: Optional[Sequence[torch.Tensor]] = None) -> None:
        """After add/prune changed P (gaussian_model.py:519-583) the bucket is re-carved (and the peer
        exchange, sized for P, replaced)."""
        if exchange is not None:
            self.exchange = exchange
        if exchanged is not None:
            self._exchanged_ids = {id(p) for p in exchanged}
        self._routed = [p for p in params if id(p) in self._exchanged_ids]
        rest = [p for p in params if id(p) not in self._exchanged_ids]
        self.bucket = FusedGradBucket(rest) if rest else None
        if optimizer is not None:
            self.optimizer = optimizer

    def resize(self, keep: Optional[torch.Tensor] = None, new: Optional[dict] = None):
        """Prune (``keep`` mask) and / or grow (``new`` raw parameters) the submap held by the fused optimiser:
        the reference's ``prune_points`` (mapper.py:206-207, 222-223) and ``add_points`` (mapper.py:255)
        with the optimiser state carried over (:meth:`eags_raster.peer.PeerAdam.resized`).  Collective; the
        exchange and the fused optimiser are replaced (returned as ``(exchange, fused_optimizer)``); a
        ``loss_fn`` that closed over the old ``fused_optimizer`` must read ``mapper.fused_optimizer`` instead."""
        if self.fused_optimizer is None:
            raise ValueError("ShardedMapper.resize: needs fused_optimizer= (torch parameters are resized by the "
                             "caller, followed by rebuild())")
        self.exchange.check()
        self.fused_optimizer = self.fused_optimizer.resized(keep=keep, new=new)
        self.exchange = self.fused_optimizer.ex
        return self.exchange, self.fused_optimizer

    def step(self, views: Sequence, loss_fn: Callable, reg_fn: Optional[Callable] = None) -> Dict:
        """One optimisation step over ``views`` (all ranks pass the same list).  Returns the local
        loss sum as a *tensor* (no host sync) and the number of views rendered locally."""
        world, rank = _world(), _rank()
        mine = [views[i] for i in shard_views(len(views), rank, world)]
        if self.bucket is not None:
            self.bucket.attach()
        local = None

        def render_all():
            nonlocal local
            for v in mine:
                loss = loss_fn(v)
                loss.backward()
                local = loss.detach() if local is None else local + loss.detach()

        if self.exchange is not None:
            if not mine:
                raise ValueError(f"ShardedMapper(exchange=...): rank {rank} has no view among {len(views)}; "
                                 "every rank must render at least one per step")
            for p in self._routed:
                p.grad = None        # the routed pass returns the all-rank result; nothing to add it to
            with self.exchange.route(scale=1.0 / len(views), local_views=len(mine),
                                     optimizer=self.fused_optimizer):
                render_all()
        else:
            render_all()
        if self.bucket is not None:
            self.bucket.all_reduce(divide_by=float(len(views)))
        if reg_fn is not None:
            reg_fn().backward()      # accumulates into the (already averaged) gradients, same on all ranks
        if self.optimizer is not None:
            self.optimizer.step()
        self._steps += 1
        if self.exchange is not None and self.check_every > 0 and self._steps % self.check_every == 0:
            self.exchange.check()    # raises if any exchange since the last check lost a peer
        return {"local_loss_sum": local, "local_views": len(mine)}
