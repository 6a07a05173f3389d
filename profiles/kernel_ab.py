"""Per-kernel device times (CUDA events recorded by the library around every launcher, L2 flushed before
each forward and each backward) for combinations of the runtime options -- one process, one GPU.

usage: python profiles/kernel_ab.py [--workloads c2,c3,c4] [--opts preprocess_agg=0/1,radix_sort=0/1]
                                    [--orders random,tile] [--reps 20]
Every printed line is one (workload, storage order, option setting): us per kernel, forward, backward.
"""
import argparse
import itertools
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402

from eags_raster import _lib, rasterize as rz, scene  # noqa: E402
from gaussian_rasterizer import GaussianRasterizationSettings  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workloads", default="c2,c3,c4")
ap.add_argument("--opts", default="preprocess_agg=1,radix_sort=1")
ap.add_argument("--orders", default="random")
ap.add_argument("--reps", type=int, default=20)
ap.add_argument("--variant", type=int, default=0)
ap.add_argument("--json", default=None)
a = ap.parse_args()
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
axes = []
for item in a.opts.split(","):
    name, vals = item.split("=")
    axes.append([(name, int(v)) for v in vals.split("/")])
rows = []
for wl in a.workloads.split(","):
    cam = scene.workload_camera(wl)
    s = GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=torch.zeros(3, device=dev),
        scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(dev), projmatrix=cam.projmatrix.to(dev), sh_degree=0,
        campos=cam.campos.to(dev), prefiltered=False, debug=False)
    H, W = cam.H, cam.W
    dC = torch.sign(torch.randn(3, H, W, device=dev)) / (3 * H * W)
    dD = torch.sign(torch.randn(1, H, W, device=dev)) / (H * W)
    for order in a.orders.split(","):
        sc = scene.make_scene(wl, seed=0, order=order)
        d = {k: sc[k].to(dev) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
        for combo in itertools.product(*axes):
            for name, v in combo:
                _lib.set_option(name, v)
            per, tf, tb = {}, 0.0, 0.0
            _lib.set_option("kernel_timing", 1)
            for i in range(a.reps + 3):
                flush.zero_()
                e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
                e[0].record()
                out = rz.rasterize_forward(s, a.variant, d["means3D"], d["opacities"], shs=d["shs"], scales=d["scales"],
                                           rotations=d["rotations"], want_n_touched=bool(a.variant))
                e[1].record()
                flush.zero_()
                e[2].record()
                rz.rasterize_backward(s, out[5], d["means3D"], d["opacities"], d["shs"], None, d["scales"],
                                      d["rotations"], None, dC, dD, None, want_shs=False, want_tau=bool(a.variant))
                e[3].record()
                torch.cuda.synchronize()
                spans = _lib.kernel_times()
                if i >= 3:
                    tf += e[0].elapsed_time(e[1])
                    tb += e[2].elapsed_time(e[3])
                    for name, ms in spans:
                        per[name] = per.get(name, 0.0) + ms
            _lib.set_option("kernel_timing", 0)
            per = {k: v / a.reps * 1e3 for k, v in per.items()}
            row = {"workload": wl, "order": order, "opts": dict(combo), "K": out[5].num_rendered,
                   "fwd_us": tf / a.reps * 1e3, "bwd_us": tb / a.reps * 1e3, "us": per}
            rows.append(row)
            print(f"{wl} {order:7s} " + " ".join(f"{n}={v}" for n, v in combo) + f" | K {row['K']} | "
                  + " ".join(f"{k} {v:.1f}" for k, v in per.items()) + f" | fwd {row['fwd_us']:.1f} bwd {row['bwd_us']:.1f}",
                  flush=True)
if a.json:
    with open(a.json, "w") as f:
        json.dump(rows, f, indent=1)
