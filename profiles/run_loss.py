"""Minimal driver for ncu: the fused per-view mapping loss (forward + backward) at one image size."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402

from eags_raster.losses import mapping_losses  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--H", type=int, default=1168)
ap.add_argument("--W", type=int, default=1752)
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--time", action="store_true")
a = ap.parse_args()
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
gt = torch.rand(3, a.H, a.W, generator=g).to(dev)
img0 = (gt + 0.1 * torch.randn(3, a.H, a.W, generator=g).to(dev)).clamp(0, 1)
gt_d = (1.0 + 3.0 * torch.rand(1, a.H, a.W, generator=g)).to(dev)
dep0 = gt_d + 0.2 * torch.randn(1, a.H, a.W, generator=g).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for it in range(a.iters):
    img, dep = img0.clone().requires_grad_(True), dep0.clone().requires_grad_(True)
    flush.zero_()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    c, d = mapping_losses(img, dep, gt, gt_d, "valid_depth", 0.2)
    e[1].record()
    (c + d).backward()
    e[2].record()
    torch.cuda.synchronize()
    if a.time:
        print(f"fwd {e[0].elapsed_time(e[1]) * 1e3:.1f} us  bwd {e[1].elapsed_time(e[2]) * 1e3:.1f} us (incl. host launch gaps)")
print("ok", float(c), float(d))
