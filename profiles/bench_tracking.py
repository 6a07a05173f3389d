"""Tracking iteration (config c2: 640x480, 300 K Gaussians, pose-gradient fwd+bwd) timed three ways:
  ref-style : the reference's formulation (tracker.py:85-128) -- P-sized pose transform + autograd in torch,
              our rasterizer, torch losses
  fused-loss: same, with eags_raster.losses.tracking_losses
  fast-path : eags_raster.tracking.render_relative + tracking_losses (no P-sized torch op at all)
  fast-path+graph: the same iteration incl. the Adam step, captured once and replayed (eags_raster.graph)
usage: python profiles/bench_tracking.py [--iters 100]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200"), os.path.join(ROOT, "tests")]
import torch  # noqa: E402

import bench  # noqa: E402
from eags_raster.losses import tracking_losses  # noqa: E402
from eags_raster.tracking import render_relative  # noqa: E402
from gaussian_rasterizer import GaussianRasterizer  # noqa: E402
from test_gpu_shims import _quat_mul, _quat_to_R  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=100)
ap.add_argument("--workload", default="c2")
a = ap.parse_args()
dev = torch.device("cuda:0")
sc, cam, s = bench.make_workload(a.workload, dev, 0)
base = {k: sc[k].to(dev) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
gt_c, gt_d = sc["gt_color"].to(dev), sc["gt_depth"].to(dev)
mask = torch.ones(1, cam.H, cam.W, dtype=torch.bool, device=dev)
raster = GaussianRasterizer(raster_settings=s)


def torch_losses(color, depth, alpha):
    cl = torch.abs(color - gt_c) + 1e-8
    dl = (torch.abs(depth - gt_d) + 1e-8) * mask
    a3 = alpha ** 3
    cl = cl * a3
    dl = dl * a3
    return cl.sum() / torch.sum(cl > 0), dl.sum() / torch.sum(dl > 0)


def make_step(kind):
    cr = torch.tensor([1.0, 0.004, -0.003, 0.002], device=dev, requires_grad=True)
    ct = torch.tensor([0.01, -0.005, 0.008], device=dev, requires_grad=True)
    graph = kind == "fast-path+graph"
    opt = torch.optim.Adam([cr, ct], lr=1e-4, capturable=graph)

    def step():
        if graph:
            if cr.grad is not None:
                opt.zero_grad(set_to_none=False)
        else:
            opt.zero_grad(set_to_none=True)
        if kind.startswith("fast-path"):
            color, depth, alpha, _ = render_relative(s, base["means3D"], base["opacities"], base["shs"],
                                                     base["scales"], base["rotations"], cr, ct)
        else:
            m3 = base["means3D"] @ _quat_to_R(cr).T + ct
            rq = _quat_mul((cr / cr.norm())[None], base["rotations"])
            color, depth, alpha, _ = raster(means3D=m3, means2D=torch.zeros_like(m3, requires_grad=True),
                                            opacities=base["opacities"], colors_precomp=None, shs=base["shs"],
                                            scales=base["scales"], rotations=rq, cov3D_precomp=None)
        if kind == "ref-style":
            lc, ld = torch_losses(color, depth, alpha)
        else:
            lc, ld = tracking_losses(color, depth, alpha, gt_c, gt_d, mask, soft_alpha=True)
        (0.6 * lc + 0.4 * ld).backward()
        opt.step()
    return step


for kind in ("ref-style", "fused-loss", "fast-path", "fast-path+graph"):
    step = make_step(kind)
    for _ in range(10):
        step()
    if kind.endswith("+graph"):       # the whole tracking iteration (incl. Adam) replayed from a CUDA graph
        from eags_raster.graph import GraphedIteration
        gi = GraphedIteration(step, warmup=1)
        step = gi.replay
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    if kind.endswith("+graph"):
        gi.check()
    print(f"{kind:16s}: {ms:.4f} ms / tracking iteration  ({1e3 / ms:.0f} it/s)")
