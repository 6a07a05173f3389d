"""One mapping iteration as the reference's optimize_submap runs it (src/entities/mapper.py:143-170):
activations (exp / sigmoid / normalise) -> render -> (1 - l) L1 + l (1 - SSIM) colour loss, masked L1 depth
loss, isotropic regulariser -> backward -> Adam step; on the mapping shapes c3 / c4.
  ref-style : our rasterizer, the reference's losses in torch ops (conv2d SSIM, boolean-mask L1, isotropic)
  fused     : the fused loss kernels of eags_raster.losses (masked L1 means, SSIM, isotropic)
  graph     : `fused`, replayed from a CUDA graph (eags_raster.graph.GraphedIteration)
usage: python profiles/bench_mapping.py [--workload c3] [--iters 100]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch
import torch.nn.functional as F
import bench
from eags_raster.graph import GraphedIteration
from eags_raster.losses import isotropic_loss, masked_l1_means, ssim
from gaussian_rasterizer import GaussianRasterizer

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c3")
ap.add_argument("--iters", type=int, default=100)
a = ap.parse_args()
dev = torch.device("cuda:0")
sc, cam, settings = bench.make_workload(a.workload, dev, 0)
raster = GaussianRasterizer(raster_settings=settings)
gt_c, gt_d = sc["gt_color"].to(dev), sc["gt_depth"].to(dev)
mask = (gt_d > 0).squeeze(0)
lam = 0.2
# raw parameters as GaussianModel keeps them (gaussian_model.py: log-scale, logit-opacity, raw quaternion)
raw = {"xyz": sc["means3D"].to(dev).clone(), "scaling": sc["scales"].to(dev).log(),
       "rotation": sc["rotations"].to(dev).clone(), "opacity": torch.logit(sc["opacities"].to(dev).clamp(1e-4, 1 - 1e-4)),
       "features": sc["shs"].to(dev).clone()}
params = {k: v.requires_grad_(True) for k, v in raw.items()}
opt = torch.optim.Adam([{"params": [params["xyz"]], "lr": 1e-4}, {"params": [params["scaling"]], "lr": 5e-3},
                        {"params": [params["rotation"]], "lr": 1e-3}, {"params": [params["opacity"]], "lr": 5e-2},
                        {"params": [params["features"]], "lr": 2.5e-3}], eps=1e-15, capturable=True, fused=True)
g1 = torch.tensor([torch.exp(torch.tensor(-(x - 5) ** 2 / 4.5)) for x in range(11)])
g1 = (g1 / g1.sum()).unsqueeze(1)
win = (g1 @ g1.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().to(dev)


def torch_ssim(i1, i2):
    i1, i2 = i1[None], i2[None]
    mu1, mu2 = F.conv2d(i1, win, padding=5, groups=3), F.conv2d(i2, win, padding=5, groups=3)
    s11 = F.conv2d(i1 * i1, win, padding=5, groups=3) - mu1 * mu1
    s22 = F.conv2d(i2 * i2, win, padding=5, groups=3) - mu2 * mu2
    s12 = F.conv2d(i1 * i2, win, padding=5, groups=3) - mu1 * mu2
    return (((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s11 + s22 + 9e-4))).mean()


def iteration(fused):
    opt.zero_grad(set_to_none=False) if params["xyz"].grad is not None else None
    scales, opac = torch.exp(params["scaling"]), torch.sigmoid(params["opacity"])
    rots = F.normalize(params["rotation"])
    m2d = torch.zeros_like(params["xyz"], requires_grad=True)
    image, depth, alpha, radii = raster(means3D=params["xyz"], means2D=m2d, opacities=opac, shs=params["features"],
                                        colors_precomp=None, scales=scales, rotations=rots, cov3D_precomp=None)
    if fused:
        l1c, depth_loss = masked_l1_means(image, depth, gt_c, gt_d, mask)
        color_loss = (1 - lam) * l1c + lam * (1 - ssim(image, gt_c))
        reg = isotropic_loss(scales)
    else:
        color_loss = (1 - lam) * (image[:, mask] - gt_c[:, mask]).abs().mean() + lam * (1 - torch_ssim(image, gt_c))
        depth_loss = (depth[:, mask] - gt_d[0][mask]).abs().mean()
        ms = scales.mean(dim=1, keepdim=True)
        reg = (scales - ms * torch.ones_like(scales)).abs().mean()
    total = color_loss + depth_loss + reg
    total.backward()
    opt.step()
    return total.detach()


def timed(fn):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / a.iters


res = {"workload": f"{a.workload}: {cam.W}x{cam.H}, {raw['xyz'].shape[0]} Gaussians, mapping iteration incl. Adam step"}
res["ref_style_ms"] = timed(lambda: iteration(False))
res["fused_ms"] = timed(lambda: iteration(True))
gi = GraphedIteration(lambda: iteration(True), warmup=2)
res["graph_ms"] = timed(gi.replay)
gi.check()
res["it_per_s"] = {k[:-3]: 1e3 / v for k, v in res.items() if k.endswith("_ms")}
print(json.dumps(res))
