"""Host-side cost of one fwd+loss+bwd step through the public API: tiny cloud (GPU work negligible),
wall clock per step in each sync mode = the pace the Python side can sustain."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch
import bench
from eags_raster import rasterize as rz
from eags_raster.losses import masked_l1_means
from gaussian_rasterizer import GaussianRasterizer

dev = torch.device("cuda:0")
sc, cam, settings = bench.make_workload("c2", dev, 0, P=2000)
keys = ("means3D", "scales", "rotations", "opacities")
d = {k: sc[k].to(dev) for k in keys + ("shs",)}
gt_c, gt_d = sc["gt_color"].to(dev), sc["gt_depth"].to(dev)
raster = GaussianRasterizer(raster_settings=settings)

def step(part):
    leaf = {k: d[k].detach().requires_grad_(True) for k in keys}
    m2d = torch.zeros_like(leaf["means3D"], requires_grad=True)
    color, depth, alpha, radii = raster(means3D=leaf["means3D"], means2D=m2d, opacities=leaf["opacities"], shs=d["shs"],
                                        colors_precomp=None, scales=leaf["scales"], rotations=leaf["rotations"], cov3D_precomp=None)
    if part == "fwd":
        return
    lc, ld = masked_l1_means(color, depth, gt_c, gt_d)
    loss = lc + ld
    if part == "fwd+loss":
        return
    loss.backward()

for mode in ("optimistic", "nosync", "exact"):
    rz.set_sync_mode(mode)
    for part in ("fwd", "fwd+loss", "all"):
        for _ in range(30): step(part)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(500): step(part)
        torch.cuda.synchronize()
        print(f"{mode:10s} {part:9s} {(time.perf_counter()-t0)*1e3/500:.4f} ms/step")

if os.environ.get("PROFILE"):
    import cProfile, pstats, io
    rz.set_sync_mode("optimistic")
    pr = cProfile.Profile(); pr.enable()
    for _ in range(300): step("all")
    torch.cuda.synchronize(); pr.disable()
    st = io.StringIO(); pstats.Stats(pr, stream=st).sort_stats("tottime").print_stats(28); print(st.getvalue()[:6000])
