"""The two round-2 binning kernels against their round-1 counterparts and the oracle.

* ``preprocess_agg`` (CTA-aggregated preprocess_fwd: shared-memory tile histogram, one global atomic per
  (CTA, touched tile)) vs the one-atomic-per-key kernel;
* ``radix_sort`` = 2 (one-pass bucket sort over the tile's depth range + thread-local insertion finish, the
  default), 1 (per-tile LSD radix sort on the varying depth bits + tie fix-up by Gaussian index) vs 0, the
  bitonic network on the 64-bit (depth | index) composite;

on the exact path (count -> scan -> duplicate -> sort) and on the bucketed path, for ordinary scenes and
for single-tile lists with runs of exactly equal depth of length 2, 5, 40 (> TIE_RUN_MAX: bitonic
fallback) and n (every key equal).  Everything compared is integer / index work: bit-exact, and the
rendered images of all variants must be bit-identical to each other.
"""
import itertools

import numpy as np
import pytest
import torch

from eags_raster import _lib, rasterize as rz
from eags_raster import scene, state_array
from helpers import cuda_forward, oracle_forward

pytestmark = pytest.mark.gpu
NAMES = ("ranges", "point_list", "point_list_keys", "n_contrib", "tiles_touched")


def _run(cam, sc):
    st, d, out = cuda_forward(cam, sc)
    state = out[5]
    K = state.num_rendered
    arrays = {n: state_array(state, n) for n in NAMES}
    arrays["point_list"] = arrays["point_list"][:K]
    arrays["point_list_keys"] = arrays["point_list_keys"][:K]
    return out, arrays


def _all_variants(cam, sc, o):
    """Runs agg x radix x {exact, bucketed}; returns nothing, asserts everything."""
    key = (torch.cuda.current_device(), cam.W, cam.H)
    K, G = o["num_rendered"], o["grid"][0] * o["grid"][1]
    ref_imgs = None
    try:
        for agg, radix, bucketed in itertools.product((0, 1), (0, 1, 2), (False, True)):
            _lib.set_option("preprocess_agg", agg)
            _lib.set_option("radix_sort", radix)
            rz.set_bucketed(bucketed)
            if bucketed:
                rz._tile_cap.pop(key, None)
                _run(cam, sc)                      # learns the bucket size on the exact path
                assert rz._tile_cap[key] > 0
            out, got = _run(cam, sc)
            tag = f"agg={agg} radix={radix} bucketed={bucketed}"
            assert out[5].num_rendered == K, tag
            np.testing.assert_array_equal(got["ranges"].cpu().numpy().view(np.uint32)[:G], o["ranges"], err_msg=tag)
            np.testing.assert_array_equal(got["point_list"].cpu().numpy().view(np.uint32), o["point_list"], err_msg=tag)
            np.testing.assert_array_equal(got["point_list_keys"].cpu().numpy().view(np.uint64), o["keys"], err_msg=tag)
            np.testing.assert_array_equal(got["tiles_touched"].cpu().numpy().view(np.uint32)[: o["P"]],
                                          o["tiles_touched"], err_msg=tag)
            np.testing.assert_array_equal(out[3].cpu().numpy(), o["radii"], err_msg=tag)
            imgs = [t.clone() for t in out[:3]] + [got["n_contrib"].clone()]
            if ref_imgs is None:
                ref_imgs = imgs
            else:
                for a, b in zip(imgs, ref_imgs):
                    assert torch.equal(a, b), tag
    finally:
        _lib.set_option("preprocess_agg", 1)
        _lib.set_option("radix_sort", 2)
        rz.set_bucketed(True)


@pytest.mark.parametrize("name,P,rot", [("ragged", 3000, (2, -1)), ("c1", 50_000, (0, 0)), ("tum_crop", 120_000, (1, 2))])
def test_variants_on_scenes(name, P, rot):
    sc = scene.make_scene(name, P=P, seed=5)
    cam = scene.workload_camera(name, w2c=scene.rot_xy(*rot))
    _all_variants(cam, sc, oracle_forward(cam, sc, skip_blend=True))


def test_agg_spans_more_than_one_cta_and_image_order():
    """P large enough that every SM's CTA walks several Gaussians per thread; stored in tile order (whole runs
    of a bucket reserved by one atomic) and in random order."""
    for order in ("random", "tile", "raster"):
        sc = scene.make_scene("c2", P=400_000, seed=1, order=order)
        cam = scene.workload_camera("c2")
        o = oracle_forward(cam, sc, skip_blend=True)
        key = (torch.cuda.current_device(), cam.W, cam.H)
        rz._tile_cap.pop(key, None)
        _run(cam, sc)
        out, got = _run(cam, sc)
        K = o["num_rendered"]
        assert out[5].num_rendered == K
        np.testing.assert_array_equal(got["point_list"].cpu().numpy().view(np.uint32), o["point_list"], err_msg=order)
        np.testing.assert_array_equal(got["point_list_keys"].cpu().numpy().view(np.uint64), o["keys"], err_msg=order)


def test_clustered_depths_fall_back_to_the_radix_sort():
    """Two thin surfaces (depth spread 1e-3 of the tile's range each): most keys of a tile land in a few of the
    bucket sort's sub-ranges, so the tile is handed to the radix sort -- same result."""
    g = torch.Generator().manual_seed(9)
    sc = scene.make_scene("c1", P=60_000, seed=9)
    z_new = torch.where(torch.rand(60_000, generator=g) < 0.5, 1.5, 3.5) + 1e-3 * torch.rand(60_000, generator=g)
    ratio = (z_new / sc["means3D"][:, 2].abs().clamp_min(1e-6))
    front = sc["means3D"][:, 2] > 0
    sc["means3D"] = torch.where(front[:, None], sc["means3D"] * ratio[:, None], sc["means3D"]).contiguous()
    cam = scene.workload_camera("c1")
    _all_variants(cam, sc, oracle_forward(cam, sc, skip_blend=True))


def _tie_scene(P, run, seed=0):
    """All Gaussians inside one 16x16 tile; depths repeat in runs of `run` exactly equal values, stored so
    that equal depths are far apart in memory (the bucket order is arbitrary anyway)."""
    g = torch.Generator().manual_seed(seed)
    cam = scene.workload_camera("tiny")
    u = 20.0 + 8.0 * torch.rand(P, generator=g)
    v = 20.0 + 8.0 * torch.rand(P, generator=g)
    n_distinct = max(1, (P + run - 1) // run)
    zs = 1.0 + 3.0 * torch.rand(n_distinct, generator=g)
    z = zs[torch.arange(P) % n_distinct]
    x = (u + 0.5 - cam.cx) / cam.fx * z
    y = (v + 0.5 - cam.cy) / cam.fy * z
    q = torch.randn(P, 4, generator=g)
    sc = dict(means3D=torch.stack([x, y, z], 1), scales=(z / cam.fx)[:, None] * (0.2 + 0.3 * torch.rand(P, 3, generator=g)),
              rotations=q / q.norm(dim=1, keepdim=True), opacities=torch.full((P, 1), 0.02),
              shs=torch.rand(P, 1, 3, generator=g))
    return cam, {k: t.float().contiguous() for k, t in sc.items()}


@pytest.mark.parametrize("P,run", [(7, 7), (100, 2), (1000, 5), (2048, 40), (2049, 3), (6000, 5), (6000, 6000),
                                   (16384, 33), (17000, 2)])
def test_depth_tie_runs(P, run):
    cam, sc = _tie_scene(P, run)
    o = oracle_forward(cam, sc, skip_blend=True)
    # the view transform must keep the ties exact: z is copied, the camera is the identity
    assert len(np.unique(o["depths"][o["radii"] > 0])) <= (P + run - 1) // run
    _all_variants(cam, sc, o)


def test_tile_order_changes_no_rendered_value():
    """eags_raster.reorder.tile_order: the same Gaussians stored by projected tile render the same image
    bit for bit (the blend order inside a tile is (depth, index); no exact depth ties in this scene), and the
    radii / gradients are the permuted originals."""
    from eags_raster import reorder
    from gaussian_rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda")
    sc = scene.make_scene("c1", P=80_000, seed=13)
    cam = scene.workload_camera("c1", w2c=scene.rot_xy(2, 3))
    s = GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=torch.zeros(3, device=dev),
        scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(dev), projmatrix=cam.projmatrix.to(dev), sh_degree=0,
        campos=cam.campos.to(dev), prefiltered=False, debug=False)
    keys = ("means3D", "scales", "rotations", "opacities", "shs")
    a = {k: sc[k].to(dev) for k in keys}
    perm = reorder.tile_order(a["means3D"], s)
    assert sorted(perm.tolist()) == list(range(80_000))
    b = dict(zip(keys, reorder.apply_order(perm, [a[k] for k in keys])))

    def render(t):
        leaf = {k: v.clone().requires_grad_(True) for k, v in t.items()}
        color, depth, alpha, radii = GaussianRasterizer(raster_settings=s)(
            means3D=leaf["means3D"], means2D=torch.zeros_like(leaf["means3D"], requires_grad=True),
            opacities=leaf["opacities"], shs=leaf["shs"], colors_precomp=None, scales=leaf["scales"],
            rotations=leaf["rotations"], cov3D_precomp=None)
        (color.sum() + depth.sum()).backward()
        return color, depth, alpha, radii, leaf["means3D"].grad

    ca, da, aa, ra, ga = render(a)
    cb, db, ab, rb, gb = render(b)
    assert torch.equal(ca, cb) and torch.equal(da, db) and torch.equal(aa, ab)
    assert torch.equal(ra[perm], rb)
    assert float((ga[perm] - gb).abs().max()) <= 1e-5 * float(ga.abs().max())
