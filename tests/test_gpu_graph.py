"""CUDA-graph replay of a whole iteration built from the public API (eags_raster/graph.py): replays give
the eager results (to the run-to-run rounding of blend_bwd's atomics), follow in-place input updates,
and an input that outgrows the captured buffers is reported, not rendered silently."""
import pytest
import torch

from eags_raster import scene
from eags_raster.graph import GraphedIteration
from eags_raster.losses import masked_l1_means
from gaussian_rasterizer import GaussianRasterizationSettings, GaussianRasterizer

pytestmark = pytest.mark.gpu
KEYS = ("means3D", "scales", "rotations", "opacities")


def _setup(P=20_000, name="c1"):
    dev = torch.device("cuda")
    sc = scene.make_scene(name, P=P, seed=3)
    cam = scene.workload_camera(name)
    s = GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(dev),
        projmatrix=cam.projmatrix.to(dev), sh_degree=0, campos=cam.campos.to(dev), prefiltered=False, debug=False)
    leaf = {k: sc[k].to(dev).requires_grad_(True) for k in KEYS}
    shs, gt_c, gt_d = sc["shs"].to(dev), sc["gt_color"].to(dev), sc["gt_depth"].to(dev)
    raster = GaussianRasterizer(raster_settings=s)

    def fn():
        for t in leaf.values():
            t.grad = None
        m2d = torch.zeros_like(leaf["means3D"], requires_grad=True)
        color, depth, alpha, radii = raster(means3D=leaf["means3D"], means2D=m2d, opacities=leaf["opacities"], shs=shs,
                                            colors_precomp=None, scales=leaf["scales"], rotations=leaf["rotations"],
                                            cov3D_precomp=None)
        lc, ld = masked_l1_means(color, depth, gt_c, gt_d)
        loss = lc + ld
        loss.backward()
        return loss.detach(), {k: leaf[k].grad for k in KEYS}

    return leaf, fn


def test_replay_matches_eager_and_follows_inplace_updates():
    leaf, fn = _setup()
    gi = GraphedIteration(fn)
    for trial in range(2):
        loss_e, grads_e = fn()
        loss_e, grads_e = loss_e.clone(), {k: v.clone() for k, v in grads_e.items()}
        loss_g, grads_g = gi.replay()
        gi.check()
        assert abs(float(loss_g) - float(loss_e)) <= 1e-6 * abs(float(loss_e))
        for k in KEYS:
            ref = grads_e[k]
            assert float((grads_g[k] - ref).abs().max()) <= 1e-5 * float(ref.abs().max()), k
        with torch.no_grad():                       # an optimiser step: parameters change in place
            leaf["means3D"].add_(0.01 * torch.randn_like(leaf["means3D"]))
            leaf["opacities"].mul_(0.9)


def test_outgrown_capture_is_reported():
    from eags_raster import rasterize as rz
    rz._capacity.clear()      # forget the high-water marks earlier tests left behind
    rz._tile_cap.clear()
    leaf, fn = _setup(P=5_000)
    gi = GraphedIteration(fn)
    gi.replay()
    gi.check()
    with torch.no_grad():
        leaf["scales"].mul_(6.0)                    # every splat now covers ~36x more tiles
    gi.replay()
    with pytest.raises(RuntimeError, match="not rendered"):
        gi.check()
    gi2 = GraphedIteration(fn)                       # re-capture with the raised capacities works
    gi2.replay()
    gi2.check()


def test_overflow_of_an_earlier_replay_is_still_reported():
    """check() once per many replays.  An iteration that outgrew the capture in the MIDDLE of a run
    (and was not rendered) must be reported even though the last replay fitted again: the status block the
    replays copy out is a sticky device-side maximum, not the last replay's."""
    from eags_raster import rasterize as rz
    rz._capacity.clear()
    rz._tile_cap.clear()
    leaf, fn = _setup(P=5_000)
    gi = GraphedIteration(fn)
    gi.replay()
    gi.check()
    with torch.no_grad():
        leaf["scales"].mul_(6.0)
    gi.replay()                                      # outgrows the capture: not rendered
    with torch.no_grad():
        leaf["scales"].mul_(1.0 / 6.0)
    for _ in range(3):
        gi.replay()                                  # fits again
    with pytest.raises(RuntimeError, match="not rendered"):
        gi.check()


def test_camera_views_are_recopied_by_every_replay():
    """the reference passes w2c.transpose(0, 1) (a non-contiguous view, utils.py:97).  Its contiguous
    copy must be a node of the graph -- an in-place camera update between replays has to show."""
    dev = torch.device("cuda")
    sc = scene.make_scene("c1", P=10_000, seed=3)
    cam = scene.workload_camera("c1")
    w2c = cam.w2c.to(dev).clone()
    proj_t = cam.projmatrix_raw.to(dev)
    view_t = w2c.transpose(0, 1)                     # non-contiguous, shares storage with w2c
    full = torch.empty(4, 4, device=dev)
    campos = torch.empty(3, device=dev)
    d = {k: sc[k].to(dev) for k in KEYS + ("shs",)}

    def render():
        full.copy_(view_t @ proj_t)
        campos.copy_(-(w2c[:3, :3].t() @ w2c[:3, 3]))
        s = GaussianRasterizationSettings(
            image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
            bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=view_t, projmatrix=full, sh_degree=0,
            campos=campos, prefiltered=False, debug=False)
        m2d = torch.zeros_like(d["means3D"], requires_grad=True)
        color, depth, alpha, radii = GaussianRasterizer(raster_settings=s)(
            means3D=d["means3D"], means2D=m2d, opacities=d["opacities"], shs=d["shs"], colors_precomp=None,
            scales=d["scales"], rotations=d["rotations"], cov3D_precomp=None)
        return color

    gi = GraphedIteration(render)
    a = gi.replay().clone()
    with torch.no_grad():
        w2c.copy_(scene.rot_xy(3, -2).to(dev))      # the camera moves, in place
    b = gi.replay().clone()
    gi.check()
    ref = render()
    assert not torch.equal(a, b)
    assert float((b - ref).abs().max()) <= 1e-5
