"""Host logic of sharded mapping (eags_raster.sharded) on CPU with gloo, world_size 2.

The render function is a small differentiable stand-in (no rasterizer on CPU -- the CUDA path has no
CPU fallback); what is under test is the N>1 plumbing: view sharding, the fused gradient bucket,
ONE all-reduce per step, identical optimiser steps on every rank, consistent prune masks and the
collective-free submap placement for parallel loop closure.
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_params(P=257, seed=3):
    g = torch.Generator().manual_seed(seed)
    xyz = torch.randn(P, 3, generator=g).requires_grad_(True)
    scale = torch.randn(P, 3, generator=g).requires_grad_(True)
    rot = torch.randn(P, 4, generator=g).requires_grad_(True)
    opa = torch.randn(P, 1, generator=g).requires_grad_(True)
    return [xyz, scale, rot, opa]


def _loss(params, view):
    """Stand-in for render + L1: depends on every parameter through their activations and on the view."""
    xyz, scale, rot, opa = params
    w = torch.tensor([1.0 + view, 0.5 - 0.1 * view, 0.25 * view + 0.1])
    q = rot / rot.norm(dim=1, keepdim=True)
    img = torch.sigmoid(opa) * (xyz * w).sum(1, keepdim=True) + torch.exp(0.1 * scale).sum(1, keepdim=True) * q[:, :1]
    return (img - 0.3 * view).abs().mean()


def _reg(params):
    s = torch.exp(0.1 * params[1])
    return 0.01 * (s - s.mean(dim=1, keepdim=True)).abs().mean()   # isotropic regulariser, losses.py


def _worker(rank, world, port, n_views, steps, out):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from eags_raster import sharded
    try:
        # replicas start different on purpose: broadcast must equalise them
        params = _make_params(seed=3 + 10 * rank)
        opt = torch.optim.SGD(params, lr=0.5, momentum=0.9)  # (Adam would amplify 1-ulp differences of ~0 gradients)
        mapper = sharded.ShardedMapper(params, opt)
        seen = []
        n_collectives = 0
        orig = dist.all_reduce

        def counting(*a, **k):
            nonlocal n_collectives
            n_collectives += 1
            return orig(*a, **k)

        dist.all_reduce = counting
        for it in range(steps):
            info = mapper.step(list(range(n_views)), lambda v: (seen.append(v), _loss(params, v))[1],
                               reg_fn=lambda: _reg(params))
            assert info["local_views"] == len(sharded.shard_views(n_views, rank, world))
        dist.all_reduce = orig
        assert n_collectives == steps, "exactly one gradient all-reduce per step"
        # p.grad are views of the fused bucket (no pack / unpack)
        base = mapper.bucket.flat.data_ptr()
        assert all(p.grad.data_ptr() >= base and p.grad.data_ptr() < base + mapper.bucket.numel * 4
                   for p in params)
        assert mapper.bucket.numel == 257 * 11
        mask = torch.zeros(257, dtype=torch.bool)
        mask[: 100 + 50 * rank] = True
        cm = sharded.consistent_mask(mask)
        torch.save({"params": [p.detach().clone() for p in params], "seen": seen, "mask": cm,
                    "submaps": sharded.assign_submaps(16)}, out % rank)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_views", [5, 8])
def test_sharded_mapping_matches_single_process(tmp_path, n_views):
    world, steps = 2, 4
    out = str(tmp_path / "r%d.pt")
    mp.spawn(_worker, args=(world, _free_port(), n_views, steps, out), nprocs=world, join=True)
    r = [torch.load(out % i) for i in range(world)]
    # every view rendered exactly once per step, split round-robin
    assert sorted(r[0]["seen"] + r[1]["seen"]) == sorted(list(range(n_views)) * steps)
    assert set(r[0]["seen"]) == set(range(0, n_views, 2)) and set(r[1]["seen"]) == set(range(1, n_views, 2))
    # replicas stay bit-identical
    for a, b in zip(r[0]["params"], r[1]["params"]):
        assert torch.equal(a, b)
    # and equal the single-process optimisation over all views (mean of view losses + regulariser)
    sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
    params = _make_params(seed=3)
    opt = torch.optim.SGD(params, lr=0.5, momentum=0.9)  # (Adam would amplify 1-ulp differences of ~0 gradients)
    for it in range(steps):
        opt.zero_grad(set_to_none=True)
        loss = sum(_loss(params, v) for v in range(n_views)) / n_views + _reg(params)
        loss.backward()
        opt.step()
    for a, b in zip(r[0]["params"], params):
        assert torch.allclose(a, b.detach(), rtol=1e-5, atol=1e-6)
    # prune masks agree (AND over ranks); submaps are placed without overlap
    assert torch.equal(r[0]["mask"], r[1]["mask"]) and int(r[0]["mask"].sum()) == 100
    assert sorted(r[0]["submaps"] + r[1]["submaps"]) == list(range(16))
    assert r[0]["submaps"] == list(range(0, 16, 2))


def test_single_process_step_is_the_reference_step():
    """world_size 1, one view per step == plain zero_grad / backward / step (mapper.py:152-170)."""
    sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
    from eags_raster import sharded
    pa, pb = _make_params(), _make_params()
    oa, ob = torch.optim.Adam(pa, lr=1e-2), torch.optim.Adam(pb, lr=1e-2)
    m = sharded.ShardedMapper(pa, oa)
    for it in range(3):
        m.step([it], lambda v: _loss(pa, v) + _reg(pa))
        ob.zero_grad(set_to_none=True)
        (_loss(pb, it) + _reg(pb)).backward()
        ob.step()
    for a, b in zip(pa, pb):
        assert torch.equal(a, b)
    assert sharded.shard_views(8, 3, 4) == [3, 7] and sharded.shard_views(3, 3, 4) == []
