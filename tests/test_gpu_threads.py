"""Re-entrancy: the reference renders from two Python threads at once (main thread: tracking / mapping,
loop-closure thread: registration, src/entities/lc.py:60,625).  Two threads hammer the operator on their
own CUDA streams with different scenes and image sizes; every result must equal the single-threaded
one (forward bit-exact, gradients to the atomics' rounding)."""
import threading

import pytest
import torch

from eags_raster import scene
from gaussian_rasterizer import GaussianRasterizationSettings, GaussianRasterizer

pytestmark = pytest.mark.gpu
KEYS = ("means3D", "scales", "rotations", "opacities")


def _job(name, P, seed):
    dev = torch.device("cuda")
    sc = scene.make_scene(name, P=P, seed=seed)
    cam = scene.workload_camera(name)
    s = GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(dev),
        projmatrix=cam.projmatrix.to(dev), sh_degree=0, campos=cam.campos.to(dev), prefiltered=False, debug=False)
    d = {k: sc[k].to(dev) for k in KEYS + ("shs", "gt_color", "gt_depth")}
    raster = GaussianRasterizer(raster_settings=s)

    def run():
        leaf = {k: d[k].detach().requires_grad_(True) for k in KEYS}
        m2d = torch.zeros_like(leaf["means3D"], requires_grad=True)
        color, depth, alpha, radii = raster(means3D=leaf["means3D"], means2D=m2d, opacities=leaf["opacities"],
                                            shs=d["shs"], colors_precomp=None, scales=leaf["scales"],
                                            rotations=leaf["rotations"], cov3D_precomp=None)
        ((color - d["gt_color"]).abs().mean() + (depth - d["gt_depth"]).abs().mean()).backward()
        return color.detach().clone(), depth.detach().clone(), radii.clone(), leaf["means3D"].grad.clone()

    return run


def test_two_threads_two_streams_same_results():
    jobs = [_job("ragged", 4000, 1), _job("c1", 30_000, 2)]
    want = [j() for j in jobs]
    torch.cuda.synchronize()
    errors = []

    def worker(i):
        try:
            stream = torch.cuda.Stream()
            with torch.cuda.stream(stream):
                for _ in range(25):
                    c, dpt, r, g = jobs[i]()
                    stream.synchronize()
                    assert torch.equal(c, want[i][0]) and torch.equal(dpt, want[i][1]) and torch.equal(r, want[i][2])
                    ref = want[i][3]
                    assert float((g - ref).abs().max()) <= 1e-5 * float(ref.abs().max())
        except BaseException as e:   # noqa: BLE001 - reported in the main thread
            errors.append((i, repr(e)))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
