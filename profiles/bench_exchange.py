"""torchrun --nproc-per-node N profiles/bench_exchange.py [--workload c2]
The gradient exchange of sharded mapping in isolation, device-timed (max over ranks, ranks aligned by a
barrier before every sample): (a) one NCCL all-reduce of the fused 11-floats/Gaussian buffer + the
1/world scale, (b) the peer_reduce kernel (barrier + rank-order sum + scale + push + barrier; its
reduce-scatter half rides inside preprocess_bwd), (c)/(d) the whole backward with each."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from eags_raster import rasterize as rz  # noqa: E402
from eags_raster.peer import PeerGradExchange  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c2")
ap.add_argument("--reps", type=int, default=50)
a = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
sc, cam, settings = bench.make_workload(a.workload, dev, seed=0, view_id=rank)
keys = ("means3D", "scales", "rotations", "opacities")
d = {k: sc[k].to(dev) for k in keys + ("shs",)}
P = d["means3D"].shape[0]
H, W = cam.H, cam.W
dC = torch.sign(torch.randn(3, H, W, device=dev)) / (3 * H * W)
dD = torch.sign(torch.randn(1, H, W, device=dev)) / (H * W)
ex = PeerGradExchange(P, dev)
flat = torch.zeros(11 * P, device=dev)
sizes = {"means3D": 3, "scales": 3, "rotations": 4, "opacities": 1}
views, off = {}, 0
for k in keys:
    views[k] = flat[off:off + P * sizes[k]].view(P, sizes[k]); off += P * sizes[k]
st = rz.rasterize_forward(settings, 0, d["means3D"], d["opacities"], shs=d["shs"], scales=d["scales"], rotations=d["rotations"])[5]
bargs = (settings, st, d["means3D"], d["opacities"], d["shs"], None, d["scales"], d["rotations"], None, dC, dD, None)


def bwd_nccl():
    with rz.grad_arena(**views):
        rz.rasterize_backward(*bargs, want_shs=False)
    dist.all_reduce(flat)
    flat.mul_(1.0 / world)


def bwd_peer():
    with ex.route(scale=1.0 / world):
        rz.rasterize_backward(*bargs, want_shs=False)


def bwd_plain():
    with rz.grad_arena(**views):
        rz.rasterize_backward(*bargs, want_shs=False)


def only_nccl():
    dist.all_reduce(flat)
    flat.mul_(1.0 / world)


def only_peer():
    ex.reduce(1.0 / world)


def timed(fn):
    for _ in range(5):
        fn()
    tot = 0.0
    for _ in range(a.reps):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    t = torch.tensor([tot / a.reps], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


res = {"world": world, "workload": a.workload, "P": P, "bytes": 44 * P}
for name, fn in [("backward_only_ms", bwd_plain), ("nccl_allreduce_ms", only_nccl), ("peer_reduce_ms", only_peer),
                 ("backward+nccl_ms", bwd_nccl), ("backward+peer_ms", bwd_peer)]:
    res[name] = timed(fn)
res["nccl_busbw_GBps"] = 2 * (world - 1) / world * 44 * P / (res["nccl_allreduce_ms"] * 1e-3) / 1e9
res["peer_busbw_GBps"] = 2 * (world - 1) / world * 44 * P / (res["peer_reduce_ms"] * 1e-3) / 1e9
assert ex.error_flag() == 0
if rank == 0:
    print(json.dumps(res))
dist.barrier()
dist.destroy_process_group()
