"""CPU checks of host-side logic added around the kernels: peer-exchange ownership / layout arithmetic
(eags_raster/peer.py must agree with csrc/collective.cu + preprocess.cu), bucketed binning sizes through
the C ABI (no compute call), capture bookkeeping of eags_raster/graph.py."""
import ctypes as C

import pytest
import torch

from eags_raster import _lib, peer
from eags_raster.graph import _CaptureCtx


@pytest.mark.parametrize("P,world", [(1, 1), (4001, 2), (1000, 3), (50_000, 8), (299_999, 8), (7, 8)])
def test_owner_chunks_partition_the_cloud_on_vector_boundaries(P, world):
    chunk, stride = peer.owner_chunk(P, world), peer.stride_of(P)
    assert chunk % 4 == 0 and stride % 4 == 0 and stride >= P and chunk * world >= P
    owners = [min(i // chunk, world - 1) for i in range(P)]
    assert owners == sorted(owners) and owners[0] == 0 and owners[-1] <= world - 1
    # a 16-byte vector of the xyz / scale / opacity regions never straddles two owners
    for k in (3, 1):
        for r in range(1, world):
            assert (k * min(P, r * chunk)) % 4 == 0 or min(P, r * chunk) == P


def test_route_struct_matches_header_layout():
    r = peer.EagsPeerRoute()
    assert C.sizeof(r) == 4 + 4 + 8 + 3 * 8 * peer.MAX_PEERS + 4 * 8
    assert peer.EagsPeerRoute.recv.offset == 16 and peer.EagsPeerRoute.flags.offset == 16 + 16 * peer.MAX_PEERS
    assert peer.EagsPeerRoute.accum.offset == 16 + 24 * peer.MAX_PEERS
    assert all(not r.accum[k] for k in range(4))      # optional: zero unless route(accum= / local_views>1)


def test_bucketed_binning_bytes():
    lib = _lib.load()
    base = lib.eags_binning_state_bytes(100_000, 0)
    assert lib.eags_binning_state_bytes_bucketed(100_000, 0, 640, 480, 0) == base
    # 1200 tiles x 1024 slots x 8 B of buckets replace the 100 000 x 8 B key list
    b = lib.eags_binning_state_bytes_bucketed(100_000, 0, 640, 480, 1024)
    assert b - base == 1200 * 1024 * 8 - 100_000 * 8
    assert lib.eags_binning_state_bytes_bucketed(100_000, 1, 640, 480, 1024) - b >= 100_000 * 8
    assert lib.eags_peer_flags_bytes() % 256 == 0 and lib.eags_peer_flags_bytes() >= (2 * peer.MAX_PEERS + 3) * 4


def test_peer_exchange_needs_cuda():
    with pytest.raises(RuntimeError, match="CUDA"):
        peer.PeerGradExchange(10, "cpu", emulate_world=2)


def test_capture_ctx_hands_out_each_status_block_once():
    if not torch.cuda.is_available():
        pytest.skip("pinned memory needs a CUDA runtime")
    ctx = _CaptureCtx(2)
    a, b = ctx.take_status(10), ctx.take_status(20)
    assert a.data_ptr() != b.data_ptr() and [c for _, c in ctx.used] == [10, 20]
    with pytest.raises(RuntimeError, match="max_renders"):
        ctx.take_status(30)


def test_sharded_mapper_splits_parameters_between_exchange_and_bucket():
    """Host logic of ShardedMapper(exchange=...): routed parameters stay out of the NCCL bucket, get their
    gradient from the routed pass only, and the route is opened for exactly the rank's views (a stand-in
    exchange records the calls; the real one needs a GPU: tests/test_gpu_peer.py)."""
    from eags_raster import sharded

    class FakeRoute:
        def __init__(self, log, **kw):
            self.log, self.kw = log, kw

        def __enter__(self):
            self.log.append(("enter", self.kw))

        def __exit__(self, *exc):
            self.log.append(("exit", None))
            return False

    class FakeExchange:
        def __init__(self):
            self.log = []

        def route(self, **kw):
            return FakeRoute(self.log, **kw)

    xyz = torch.zeros(5, 3, requires_grad=True)
    feat = torch.ones(5, 3, requires_grad=True)
    xyz.grad = torch.full((5, 3), 7.0)           # stale gradient of an earlier step
    opt = torch.optim.SGD([xyz, feat], lr=0.5)
    ex = FakeExchange()
    m = sharded.ShardedMapper([xyz, feat], opt, broadcast=False, exchange=ex, exchanged=[xyz])
    assert [id(p) for p in m.bucket.params] == [id(feat)]
    seen = []

    def loss_fn(v):
        seen.append(v)
        assert xyz.grad is None or not torch.any(xyz.grad == 7.0)      # cleared before the first pass
        return (xyz.sum() + feat.sum()) * float(v)

    out = m.step([1.0, 2.0, 3.0], loss_fn)
    assert seen == [1.0, 2.0, 3.0] and out["local_views"] == 3
    assert ex.log[0] == ("enter", {"scale": 1.0 / 3, "local_views": 3, "optimizer": None}) and ex.log[-1][0] == "exit"
    # the bucket parameter carries the mean over views (world 1): (1 + 2 + 3) / 3 = 2 -> 1 - 0.5 * 2 = 0
    assert torch.allclose(feat.detach(), torch.zeros(5, 3))
    with pytest.raises(ValueError, match="needs exchange"):
        sharded.ShardedMapper([xyz], opt, broadcast=False, exchanged=[xyz])
    # all parameters routed: no bucket, no NCCL call
    m2 = sharded.ShardedMapper([xyz], torch.optim.SGD([xyz], lr=0.1), broadcast=False, exchange=ex, exchanged=[xyz])
    assert m2.bucket is None
    m2.step([1.0], lambda v: xyz.sum())
