"""CUDA-graph replay of a fixed-shape iteration (forward -> loss -> backward) built from the public API.

The tracking loop runs 60-300 iterations per frame on tensors whose shapes do not change
(src/entities/tracker.py:181-348), and a mapping window runs 100+ iterations per submap size
(mapper.py:143-170).  Issued eagerly, one iteration costs ~0.4 ms of Python / autograd / launch time on
the host and one host wait (the key count K), which is what bounds the step once the kernels are fast.
:class:`GraphedIteration` runs ``fn`` a few times eagerly (learning the binning capacities), captures
it ONCE into a CUDA graph and replays it: the same kernels, one launch per iteration.

Safety: a replay cannot grow a buffer.  Every captured forward folds its status (K, per-tile maximum,
overflow flag) into a persistent device block (atomicMax / atomicOr: sticky over ALL replays) and copies
that block to pinned memory; :meth:`GraphedIteration.check` (call it whenever the host synchronises
anyway, e.g. once per frame) raises if ANY replay since the capture outgrew the captured capacity -- its
outputs were not rendered -- after raising the capacities so that a re-capture succeeds.
"""
from __future__ import annotations

from typing import Callable, List, Tuple

import torch

from . import rasterize as rz


class _CaptureCtx:
    """Per captured forward: a pinned status block the replay copies into, and a persistent DEVICE block
    (zeroed here, once, never by the forward) that accumulates K / largest tile with atomicMax and the
    overflow flag with atomicOr over every replay -- the pinned copy is a copy of THAT block, so check()
    sees an overflow of any replay since the capture, not just the last one."""

    def __init__(self, n_slots: int, device):
        self.free = [torch.zeros(8, dtype=torch.int64).pin_memory() for _ in range(n_slots)]
        self.sticky = [torch.zeros(8, dtype=torch.int64, device=device) for _ in range(n_slots)]
        self.used: List[Tuple[torch.Tensor, int]] = []
        self._keep: List[torch.Tensor] = []

    def take_status(self, cap: int, device) -> Tuple[torch.Tensor, torch.Tensor]:
        if not self.free:
            raise RuntimeError("GraphedIteration: more rasterizer calls in one iteration than max_renders")
        st, sticky = self.free.pop(), self.sticky.pop()
        if sticky.device != torch.device(device):
            raise RuntimeError("GraphedIteration: the captured rasterizer call runs on another device than the graph")
        self.used.append((st, cap))
        self._keep.append(sticky)
        return st, sticky


class GraphedIteration:
    """``fn()`` must read its inputs from, and leave its results in, tensors that persist between
    calls (parameters updated in place, ``.grad`` tensors, a loss tensor it returns).  Parameters whose
    ``.grad`` is reset to ``None`` inside ``fn`` should be fresh leaves (first used by the warm-up run
    here): autograd ties a leaf's gradient accumulation to the stream of its first use, and a leaf that
    already accumulated on the legacy default stream cannot do so inside a capture ("legacy stream depends
    on a capturing stream").  Keeping the ``.grad`` tensors (``zero_grad(set_to_none=False)``) works for
    any leaf."""

    def __init__(self, fn: Callable[[], object], warmup: int = 3, max_renders: int = 16, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._fn = fn
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):      # eager: learns K capacity and per-tile bucket size
                fn()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self._ctx = _CaptureCtx(max_renders, self.device)
        self.graph = torch.cuda.CUDAGraph()
        prev = rz._capture_ctx
        rz._capture_ctx = self._ctx
        try:
            with torch.cuda.graph(self.graph):
                self.result = fn()
        finally:
            rz._capture_ctx = prev
        self.replays = 0

    def replay(self):
        self.graph.replay()
        self.replays += 1
        return self.result

    def check(self) -> None:
        """Synchronise and verify that no replay outgrew the captured buffers."""
        torch.cuda.synchronize(self.device)
        idx = self.device.index
        for st, cap in self._ctx.used:
            K = int(st[0])
            s32 = st.view(torch.int32)
            max_tile, overflow = int(s32[7]), int(s32[8])
            if K > cap or overflow:
                rz._capacity[idx] = max(rz._capacity.get(idx, 0), rz._round_capacity(int(K * 1.5)))
                for key in list(rz._tile_cap):
                    if key[0] == idx:
                        rz._tile_cap[key] = max(rz._tile_cap[key], (int(max_tile * 1.5) + 127) & ~63)
                raise RuntimeError(
                    f"GraphedIteration: a replay produced {K} keys (largest tile {max_tile}) but the captured "
                    f"buffers held {cap}; that iteration was not rendered. Capacities raised: capture again.")
