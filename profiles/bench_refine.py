"""One iteration of the reference's global-map refinement (src/evaluation/evaluate_merged_map.py:60-140:
sh_degree 3 -> M = 16 SH coefficients per Gaussian, ALL parameters incl. the SH features optimised,
(1 - l) L1 + l (1 - SSIM) colour loss, Adam): the one caller of the rasterizer that exercises the
view-dependent SH forward and the 48-float dL/dSH rows of the backward (last row).
  eager : public API + fused SSIM/L1 kernels, launched from Python
  graph : the same iteration replayed from a CUDA graph (eags_raster.graph.GraphedIteration)
  m1    : the same shape with M = 1 (sh_degree 0) for the cost of the SH path alone
usage: python profiles/bench_refine.py [--workload c3] [--iters 100]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch
import torch.nn.functional as F
from eags_raster import scene
from eags_raster.graph import GraphedIteration
from eags_raster.losses import masked_l1_means, ssim
from gaussian_rasterizer import GaussianRasterizationSettings, GaussianRasterizer

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c3")
ap.add_argument("--iters", type=int, default=100)
a = ap.parse_args()
dev = torch.device("cuda:0")
lam = 0.2


def timed(fn):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / a.iters


def build(M, degree):
    sc = scene.make_scene(a.workload, seed=0, M=M)
    cam = scene.workload_camera(a.workload)
    settings = GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(dev),
        projmatrix=cam.projmatrix.to(dev), sh_degree=degree, campos=cam.campos.to(dev),
        prefiltered=False, debug=False)
    raster = GaussianRasterizer(raster_settings=settings)
    gt_c, gt_d = sc["gt_color"].to(dev), sc["gt_depth"].to(dev)
    raw = {"xyz": sc["means3D"].to(dev).clone(), "scaling": sc["scales"].to(dev).log(),
           "rotation": sc["rotations"].to(dev).clone(),
           "opacity": torch.logit(sc["opacities"].to(dev).clamp(1e-4, 1 - 1e-4)),
           "features": sc["shs"].to(dev).clone()}
    params = {k: v.requires_grad_(True) for k, v in raw.items()}
    opt = torch.optim.Adam([{"params": [params["xyz"]], "lr": 1e-4}, {"params": [params["scaling"]], "lr": 5e-3},
                            {"params": [params["rotation"]], "lr": 1e-3}, {"params": [params["opacity"]], "lr": 5e-2},
                            {"params": [params["features"]], "lr": 2.5e-3}], eps=1e-15, capturable=True, fused=True)

    def iteration():
        if params["xyz"].grad is not None:
            opt.zero_grad(set_to_none=False)
        scales, opac = torch.exp(params["scaling"]), torch.sigmoid(params["opacity"])
        rots = F.normalize(params["rotation"])
        m2d = torch.zeros_like(params["xyz"], requires_grad=True)
        image, depth, alpha, radii = raster(means3D=params["xyz"], means2D=m2d, opacities=opac,
                                            shs=params["features"], colors_precomp=None, scales=scales,
                                            rotations=rots, cov3D_precomp=None)
        l1c, _ = masked_l1_means(image, depth, gt_c, gt_d)
        total = (1 - lam) * l1c + lam * (1 - ssim(image, gt_c))
        total.backward()
        opt.step()
        return total.detach()

    return iteration, cam, raw["xyz"].shape[0]


res = {}
it16, cam, P = build(16, 3)
res["workload"] = f"{a.workload}: {cam.W}x{cam.H}, {P} Gaussians, refinement iteration (M=16, all params, Adam)"
res["eager_ms"] = timed(it16)
gi = GraphedIteration(it16, warmup=2)
res["graph_ms"] = timed(gi.replay)
gi.check()
del gi, it16
torch.cuda.empty_cache()
it1, _, _ = build(1, 0)
res["m1_eager_ms"] = timed(it1)
gi = GraphedIteration(it1, warmup=2)
res["m1_graph_ms"] = timed(gi.replay)
gi.check()
res["it_per_s"] = {k[:-3]: 1e3 / v for k, v in res.items() if k.endswith("_ms")}
print(json.dumps(res))
