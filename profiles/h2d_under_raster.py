"""Per-copy time of the e2e block's
host->device copy (16.8 MB, bench.py's e2e leg) while THIS repo's iteration graph replays on the compute stream,
with and without the per-step 16.8 MB device-to-device copy bench.py issues before each replay.

profiles/r1c/h2d_ab.txt shows the copy alone (idle, or under tensor-core matmuls) at a steady 0.310 ms from
4-KiB and from huge pages; profiles/r1c/e2e_h2d_variance.txt shows it at 0.31 ... 1.36 ms inside bench.py.  This
script separates "our kernels / our D2D copy disturb the DMA" from "the box's host side is busy".  Result
(profiles/r1c/h2d_under_raster.txt, last GPU seconds of round 1): 0.311 ms per copy (54 GB/s) idle, under graph
replay and under replay + D2D copy, from 4-KiB and from huge pages alike -> the GPU-side load is not the cause.

    python profiles/h2d_under_raster.py          # prints min / p10 / median / p90 / max per phase
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402

import bench  # noqa: E402
from eags_raster.graph import GraphedIteration  # noqa: E402
from eags_raster.losses import masked_l1_means  # noqa: E402
from gaussian_rasterizer import GaussianRasterizer  # noqa: E402

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
sc, cam, settings = bench.make_workload("c2", dev, seed=0)
keys = ("means3D", "scales", "rotations", "opacities")
order = keys + ("shs",)
devt = {k: sc[k].to(dev) for k in order}
sizes = [devt[k].numel() for k in order]
gt_c, gt_d = sc["gt_color"].to(dev), sc["gt_depth"].to(dev)
raster = GaussianRasterizer(raster_settings=settings)
static_flat = torch.cat([devt[k].reshape(-1) for k in order])
static, o = {}, 0
for k, n in zip(order, sizes):
    static[k] = static_flat[o:o + n].view(devt[k].shape)
    o += n
leaves = {k: static[k].requires_grad_(True) for k in keys}


def iteration():
    for t in leaves.values():
        t.grad = None
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    color, depth, alpha, radii = raster(means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"],
                                        shs=static["shs"], colors_precomp=None, scales=leaves["scales"],
                                        rotations=leaves["rotations"], cov3D_precomp=None)
    lc, ld = masked_l1_means(color, depth, gt_c, gt_d)
    loss = lc + ld
    loss.backward()
    return loss.detach()


gi = GraphedIteration(iteration, warmup=3)
for _ in range(3):
    gi.replay()
gi.check()

N = static_flat.numel()
host_a = torch.empty(N, dtype=torch.float32).pin_memory().fill_(1.0)
host_b, kind_b = bench.pinned_host_block(N)
host_b.fill_(1.0)
ring = [torch.empty(N, dtype=torch.float32, device=dev) for _ in range(3)]
cs = torch.cuda.Stream(device=dev)


def phase(label, reps, load):
    """load: None | "replay" | "replay+d2d".  Copies are issued back to back on the copy stream."""
    evs = []
    main = torch.cuda.current_stream()
    for i in range(reps):
        if load is not None:
            if load == "replay+d2d":
                with torch.no_grad():
                    static_flat.copy_(ring[2], non_blocking=True)
            gi.replay()
        with torch.cuda.stream(cs):
            for name, src, dst in (("4K", host_a, ring[0]), ("2M", host_b, ring[1])):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(cs)
                dst.copy_(src, non_blocking=True)
                e1.record(cs)
                evs.append((name, e0, e1))
    main.synchronize()
    cs.synchronize()
    for name in ("4K", "2M"):
        t = np.array([a.elapsed_time(b) for n_, a, b in evs if n_ == name])
        q = np.percentile(t, [0, 10, 50, 90, 100])
        print(f"{label:12s} {name} pages: per-copy ms min {q[0]:.3f} p10 {q[1]:.3f} median {q[2]:.3f} p90 {q[3]:.3f} "
              f"max {q[4]:.3f}  ({N * 4 / q[2] / 1e6:.1f} GB/s at the median)", flush=True)


ring[2].copy_(static_flat)
print("2M block:", kind_b)
phase("warm", 10, None)
for rnd in range(3):
    phase("idle", 100, None)
    phase("replay", 100, "replay")
    phase("replay+d2d", 100, "replay+d2d")
gi.check()
