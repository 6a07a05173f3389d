"""Minimal driver for ncu: a few fwd+bwd iterations of the c2 workload through the public API."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402

import bench  # noqa: E402
from eags_raster import _lib  # noqa: E402
from gaussian_rasterizer import GaussianRasterizer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c2")
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--tma", type=int, default=1)
ap.add_argument("--P", type=int, default=None)
ap.add_argument("--variant-b", action="store_true")
a = ap.parse_args()
_lib.set_option("tma", a.tma)
dev = torch.device("cuda:0")
sc, cam, s = bench.make_workload(a.workload, dev, 0, P=a.P)
keys = ("means3D", "scales", "rotations", "opacities")
d = {k: sc[k].to(dev) for k in keys + ("shs", "gt_color", "gt_depth")}
r = GaussianRasterizer(raster_settings=s)
for _ in range(a.iters):
    leaf = {k: d[k].detach().requires_grad_(True) for k in keys}
    m2d = torch.zeros_like(leaf["means3D"], requires_grad=True)
    color, depth, alpha, radii = r(means3D=leaf["means3D"], means2D=m2d, opacities=leaf["opacities"], shs=d["shs"],
                                   colors_precomp=None, scales=leaf["scales"], rotations=leaf["rotations"],
                                   cov3D_precomp=None)
    ((color - d["gt_color"]).abs().mean() + (depth - d["gt_depth"]).abs().mean()).backward()
torch.cuda.synchronize()
print("ok")
