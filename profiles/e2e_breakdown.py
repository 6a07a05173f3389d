"""Where the end-to-end (host-buffer) step spends its time: wall-clock ms/step of the c2 step with
(a) device inputs, no host sync; (b) device inputs + loss read back every step; (c) pinned host inputs
copied every step (pipelined) + loss read back (= bench.py's e2e); each in sync modes optimistic / nosync."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402

import bench  # noqa: E402
from eags_raster import rasterize as rz  # noqa: E402
from eags_raster.losses import masked_l1_means  # noqa: E402
from gaussian_rasterizer import GaussianRasterizer  # noqa: E402

dev = torch.device("cuda:0")
sc, cam, settings = bench.make_workload("c2", dev, 0)
keys = ("means3D", "scales", "rotations", "opacities")
host = {k: sc[k].pin_memory() for k in keys + ("shs",)}
devt = {k: sc[k].to(dev) for k in keys + ("shs",)}
gt_c, gt_d = sc["gt_color"].to(dev), sc["gt_depth"].to(dev)
raster = GaussianRasterizer(raster_settings=settings)
copy_stream = torch.cuda.Stream(device=dev)


def step(src, read):
    leaf = {k: src[k].detach().requires_grad_(True) for k in keys}
    m2d = torch.zeros_like(leaf["means3D"], requires_grad=True)
    color, depth, alpha, radii = raster(means3D=leaf["means3D"], means2D=m2d, opacities=leaf["opacities"], shs=src["shs"],
                                        colors_precomp=None, scales=leaf["scales"], rotations=leaf["rotations"], cov3D_precomp=None)
    lc, ld = masked_l1_means(color, depth, gt_c, gt_d)
    loss = lc + ld
    loss.backward()
    return float(loss.detach()) if read else None


def prefetch():
    with torch.cuda.stream(copy_stream):
        bufs = {k: host[k].to(dev, non_blocking=True) for k in host}
        ev = torch.cuda.Event(); ev.record(copy_stream)
    return bufs, ev


def loop(n, mode):
    if mode == "host":
        nxt = prefetch()
    for i in range(n):
        if mode == "host":
            bufs, ev = nxt
            if i + 1 < n:
                nxt = prefetch()
            torch.cuda.current_stream().wait_event(ev)
            step(bufs, True)
            for t_ in bufs.values():
                t_.record_stream(torch.cuda.current_stream())
        else:
            step(devt, mode == "dev+read")


for sync in ("optimistic", "nosync"):
    rz.set_sync_mode(sync)
    for mode in ("dev", "dev+read", "host"):
        loop(20, mode); torch.cuda.synchronize()
        t0 = time.perf_counter(); loop(200, mode); torch.cuda.synchronize()
        print(f"sync={sync:10s} {mode:9s} {(time.perf_counter() - t0) * 1e3 / 200:.4f} ms/step")
# pure host cost of issuing one step (GPU far behind: queue 50 steps on a busy GPU is not possible to isolate; instead time with tiny P)
