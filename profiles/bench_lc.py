"""Parallel loop closure shape (BASELINE.json configs[4]): independent ~300 K-Gaussian submaps, each
registered by pose-gradient iterations through variant B of the operator
(`diff_gaussian_rasterization`, src/gsr/renderer.py:131-142: theta / rho gradients, opacity image,
n_touched) -- no collective; under torchrun every rank takes its share of the submaps
(eags_raster.sharded.assign_submaps) and the job's throughput is the sum.
usage: [torchrun --nproc-per-node N] python profiles/bench_lc.py [--submaps 16] [--iters 50]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402
from eags_raster import scene  # noqa: E402
from eags_raster.graph import GraphedIteration  # noqa: E402
from eags_raster.losses import masked_l1_means  # noqa: E402
from eags_raster.sharded import assign_submaps  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--submaps", type=int, default=16)
ap.add_argument("--iters", type=int, default=50)
ap.add_argument("--launch", default="graph", choices=["graph", "eager"])
a = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
cam = scene.workload_camera("c2")
s = GaussianRasterizationSettings(
    image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=torch.zeros(3, device=dev),
    scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(dev), projmatrix=cam.projmatrix.to(dev),
    projmatrix_raw=cam.projmatrix_raw.to(dev), sh_degree=0, campos=cam.campos.to(dev), prefiltered=False, debug=False)
raster = GaussianRasterizer(raster_settings=s)
mine = assign_submaps(a.submaps, rank, world)
total_ms, total_it = 0.0, 0
for sid in mine:
    sc = scene.make_scene("c2", seed=sid)
    d = {k: sc[k].to(dev) for k in ("means3D", "scales", "rotations", "opacities", "shs", "gt_color", "gt_depth")}
    theta = torch.zeros(3, device=dev, requires_grad=True)
    rho = torch.zeros(3, device=dev, requires_grad=True)
    opt = torch.optim.Adam([theta, rho], lr=1e-4, capturable=True)

    def iteration():
        opt.zero_grad(set_to_none=False) if theta.grad is not None else None
        m2d = torch.zeros_like(d["means3D"])
        color, radii, depth, opacity, n_touched = raster(means3D=d["means3D"], means2D=m2d, opacities=d["opacities"],
                                                         shs=d["shs"], scales=d["scales"], rotations=d["rotations"],
                                                         theta=theta, rho=rho)
        lc, ld = masked_l1_means(color, depth, d["gt_color"], d["gt_depth"])
        (lc + ld).backward()
        opt.step()
        return lc.detach()

    for _ in range(3):
        iteration()
    run = iteration
    if a.launch == "graph":
        gi = GraphedIteration(iteration, warmup=1)
        run = gi.replay
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    if a.launch == "graph":
        gi.check()
    total_ms += e0.elapsed_time(e1)
    total_it += a.iters
    assert torch.isfinite(theta).all() and torch.isfinite(rho).all() and float(theta.grad.abs().sum() + rho.grad.abs().sum()) > 0
t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    ms = float(t)
    print(json.dumps({"workload": f"c5: {a.submaps} independent submaps x 300000 Gaussians, 640x480, variant B pose-gradient iterations",
                      "n_gpus": world, "launch": a.launch, "iterations_total": a.submaps * a.iters,
                      "value": a.submaps * a.iters / ms * 1e3, "unit": "it/s", "ms_per_iteration_per_gpu": ms / max(total_it, 1),
                      "collective": "none (replicas only)"}))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
