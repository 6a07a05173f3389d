"""Parity of the sm_100a CUDA path (through the C ABI) against the CPU oracle on seeded scenes.

Bars (BASELINE.json north_star): radii, tile ranges and sorted keys bit-exact; rendered colour /
depth within 1e-5 relative; gradients within 1e-3 relative (fp32, order of summation).  Pixels the
oracle flags as *fragile* (a hard threshold of SURVEY A.6 Q6/Q10 within a 1e-5 band) are excluded
from the float comparisons and their upstream gradients zeroed, exactly as in the oracle-vs-autograd
test.
"""
import numpy as np
import pytest
import torch

import oracle
from eags_raster import scene, state_array, rasterize_backward, _lib
from helpers import (assert_close_rel, cuda_forward, nonfragile, oracle_forward, rel_err)

pytestmark = pytest.mark.gpu

RTOL_IMG = 1e-5     # colour / depth / alpha: relative to max |ref|
RTOL_GRAD = 1e-3    # gradients: relative to max |ref|

CASES = [
    # workload, P, M, degree, variant, bg, cam rotation
    ("tiny", None, 1, 0, 0, (0, 0, 0), (0, 0)),
    ("ragged", None, 4, 1, 1, (0.1, 0.2, 0.3), (3, -2)),
    ("ragged", 3000, 16, 3, 0, (0.3, 0.1, 0.2), (-2, 4)),
    ("ragged", 2011, 9, 2, 0, (0, 0, 0), (1, -3)),      # 27-float SH rows (odd), P not a multiple of 32
    ("ragged", 1000, 16, 1, 1, (0, 0, 0), (0, 2)),      # more coefficients than the active degree uses
    ("ragged", 3000, 16, 3, 0, (0, 0, 0), (25, 10)),    # 58 % visible: sparse warps fetch SH rows one by one
    ("c1", 50_000, 1, 0, 0, (0, 0, 0), (0, 0)),
    ("c1", 20_000, 1, 0, 1, (0, 0, 0), (1, 1)),
]


def _np(t):
    return t.detach().cpu().numpy()


def _check_forward(o, out, state, variant):
    color, depth, third, radii, n_touched, _ = out
    P = o["P"]
    # ---- bit-exact integer / index / key work ----
    np.testing.assert_array_equal(_np(radii), o["radii"])
    np.testing.assert_array_equal(_np(state_array(state, "tiles_touched")).view(np.uint32)[:P], o["tiles_touched"])
    assert state.num_rendered == o["num_rendered"]
    K = o["num_rendered"]
    G = o["grid"][0] * o["grid"][1]
    np.testing.assert_array_equal(_np(state_array(state, "ranges")).view(np.uint32)[:G], o["ranges"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list")).view(np.uint32)[:K], o["point_list"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list_keys")).view(np.uint64)[:K], o["keys"])
    vis = o["radii"] > 0
    # depth keys / pixel means come from the fmad-free preprocess: bit-exact as well
    np.testing.assert_array_equal(_np(state_array(state, "xy"))[:P][vis], o["xy"][vis])
    rgbd = _np(state_array(state, "rgbd"))[:P]
    np.testing.assert_array_equal(rgbd[vis, 3], o["depths"][vis])
    np.testing.assert_allclose(rgbd[vis, :3], o["rgb"][vis], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(_np(state_array(state, "conic_opacity"))[:P][vis], o["conic_opacity"][vis],
                               rtol=1e-6, atol=1e-9)
    cl = _np(state_array(state, "clamped"))[:P][vis]
    np.testing.assert_array_equal(np.stack([(cl >> c) & 1 for c in range(3)], 1), o["clamped"][vis])
    # ---- images ----
    ok = nonfragile(o)
    # fragile pixels: 0.006 % (c1) .. 0.04 % (c2) of a full-size image; the 3 000-pixel test images may
    # hold a handful.  They differ by one blended splat (the 1/255 alpha cut or the T < 1e-4 stop falls on
    # the other side of the threshold in the other implementation's rounding).
    assert ok.mean() >= (0.999 if ok.size >= 100_000 else 0.99), f"fragile fraction {1 - ok.mean():.4%}"
    nc = _np(state_array(state, "n_contrib")).view(np.uint32)
    np.testing.assert_array_equal(nc[ok], o["n_contrib"][ok])
    assert_close_rel(_np(color)[:, ok], o["color"][:, ok], 1e-5, RTOL_IMG, "color")
    assert_close_rel(_np(depth)[0][ok], o["depth"][0][ok], 1e-5, RTOL_IMG, "depth")
    assert_close_rel(_np(third)[0][ok], o["third"][0][ok], 1e-5, RTOL_IMG, "alpha/opacity")
    assert_close_rel(_np(state_array(state, "final_T"))[ok], o["final_T"][ok], 1e-5, RTOL_IMG, "final_T")
    if variant == 1:
        nt = _np(n_touched)
        okg = o["n_touched_fragile"] == 0
        # a fragile pixel can flip any Gaussian it blends: allow those rows to differ
        frag_pix = (~ok).sum()
        diff = (nt != o["n_touched"]) & okg
        assert diff.sum() <= 4 * frag_pix, f"n_touched differs on {diff.sum()} Gaussians"


@pytest.mark.parametrize("wl,P,M,degree,variant,bg,rot", CASES)
@pytest.mark.parametrize("tma", [1, 0])
def test_forward_backward_parity(wl, P, M, degree, variant, bg, rot, tma):
    _lib.set_option("tma", tma)
    try:
        sc = scene.make_scene(wl, P=P, seed=11, M=M)
        cam = scene.workload_camera(wl, w2c=scene.rot_xy(*rot))
        o = oracle_forward(cam, sc, degree=degree, bg=bg, variant=variant)
        st, d, out = cuda_forward(cam, sc, degree=degree, bg=bg, variant=variant)
        state = out[5]
        _check_forward(o, out, state, variant)
        # ---- backward ----
        H, W = cam.H, cam.W
        g = torch.Generator().manual_seed(5)
        dC = torch.randn(3, H, W, generator=g)
        dD = torch.randn(1, H, W, generator=g)
        dA = torch.randn(1, H, W, generator=g)
        fr = torch.from_numpy(~nonfragile(o))
        dC[:, fr] = 0
        dD[:, fr] = 0
        dA[:, fr] = 0
        ref = oracle.backward(o, dC.numpy(), dD.numpy(), dA.numpy(), want_tau=True)
        got = rasterize_backward(st, state, d["means3D"], d["opacities"], d["shs"], None, d["scales"],
                                 d["rotations"], None, dC.cuda(), dD.cuda(), dA.cuda(), want_tau=True)
        assert rel_err(_np(got["means2D"])[:, :2], ref["dL_dmean2D"]) < RTOL_GRAD
        assert np.all(_np(got["means2D"])[:, 2] == 0)
        assert rel_err(_np(got["means3D"]), ref["dL_dmeans3D"]) < RTOL_GRAD
        assert rel_err(_np(got["scales"]), ref["dL_dscales"]) < RTOL_GRAD
        assert rel_err(_np(got["rotations"]), ref["dL_drots"]) < RTOL_GRAD
        assert rel_err(_np(got["opacities"]).reshape(-1), ref["dL_dopacity"]) < RTOL_GRAD
        assert rel_err(_np(got["shs"]), ref["dL_dshs"]) < RTOL_GRAD
        assert rel_err(_np(got["tau"]), ref["dL_dtau"]) < RTOL_GRAD
        # culled Gaussians get exactly zero gradient
        cul = o["radii"] == 0
        assert np.all(_np(got["means3D"])[cul] == 0) and np.all(_np(got["shs"])[cul] == 0)
        if degree > 0:
            # frozen SH features: dL/dmeans3D still carries the view-direction term of the SH colour
            frozen = rasterize_backward(st, state, d["means3D"], d["opacities"], d["shs"], None, d["scales"],
                                        d["rotations"], None, dC.cuda(), dD.cuda(), dA.cuda(), want_shs=False)
            assert frozen["shs"] is None
            assert rel_err(_np(frozen["means3D"]), ref["dL_dmeans3D"]) < RTOL_GRAD
    finally:
        _lib.set_option("tma", 1)


def test_colors_precomp_and_scale_modifier():
    sc = scene.make_scene("ragged", P=1500, seed=2)
    cam = scene.workload_camera("ragged")
    cp = torch.rand(1500, 3, generator=torch.Generator().manual_seed(1)).numpy()
    o = oracle_forward(cam, sc, colors_precomp=cp, scale_modifier=1.3)
    st, d, out = cuda_forward(cam, sc, colors_precomp=cp, scale_modifier=1.3)
    _check_forward(o, out, out[5], 0)
    H, W = cam.H, cam.W
    dC = torch.ones(3, H, W)
    dD = torch.ones(1, H, W)
    fr = torch.from_numpy(~nonfragile(o))
    dC[:, fr] = 0
    dD[:, fr] = 0
    ref = oracle.backward(o, dC.numpy(), dD.numpy(), None)
    got = rasterize_backward(st, out[5], d["means3D"], d["opacities"], None, torch.as_tensor(cp).cuda(),
                             d["scales"], d["rotations"], None, dC.cuda(), dD.cuda(), None)
    assert rel_err(_np(got["colors_precomp"]), ref["dL_dcolors_precomp"]) < RTOL_GRAD
    assert rel_err(_np(got["scales"]), ref["dL_dscales"]) < RTOL_GRAD
    assert got["shs"] is None


def test_cov3d_precomp_equals_scale_rotation_path():
    sc = scene.make_scene("ragged", P=1200, seed=4)
    cam = scene.workload_camera("ragged", w2c=scene.rot_xy(2, 2))
    o = oracle_forward(cam, sc, skip_blend=True)
    _, _, a = cuda_forward(cam, sc)
    _, _, b = cuda_forward(cam, sc, cov3D_precomp=o["cov3D"] * 1.0)
    # invisible rows of the oracle's cov3D are zero -> those Gaussians are culled either way
    # (det of the blurred 2-D covariance stays positive, but they were culled by depth/rect before)
    vis = o["radii"] > 0
    np.testing.assert_array_equal(_np(a[3])[vis], _np(b[3])[vis])


def _single_tile_scene(P, seed=0):
    """All Gaussians inside one 16x16 tile: exercises the big / global-memory sort variants."""
    g = torch.Generator().manual_seed(seed)
    cam = scene.workload_camera("tiny")
    u = 20.0 + 8.0 * torch.rand(P, generator=g)
    v = 20.0 + 8.0 * torch.rand(P, generator=g)
    z = 1.0 + 3.0 * torch.rand(P, generator=g)
    z[: P // 4] = z[P // 4: 2 * (P // 4)][: P // 4]          # many exact depth ties
    x = (u + 0.5 - cam.cx) / cam.fx * z
    y = (v + 0.5 - cam.cy) / cam.fy * z
    q = torch.randn(P, 4, generator=g)
    sc = dict(means3D=torch.stack([x, y, z], 1), scales=(z / cam.fx)[:, None] * (0.2 + 0.3 * torch.rand(P, 3, generator=g)),
              rotations=q / q.norm(dim=1, keepdim=True), opacities=torch.full((P, 1), 0.02),
              shs=torch.rand(P, 1, 3, generator=g))
    return cam, {k: t.float().contiguous() for k, t in sc.items()}


@pytest.mark.parametrize("P", [1, 2, 33, 2048, 2049, 5000, 16384, 17000])
def test_sort_variants_and_depth_ties(P):
    cam, sc = _single_tile_scene(P)
    o = oracle_forward(cam, sc, skip_blend=True)
    st, d, out = cuda_forward(cam, sc)
    state = out[5]
    K = o["num_rendered"]
    assert state.num_rendered == K
    G = o["grid"][0] * o["grid"][1]
    assert int(o["ranges"][:, 1].astype(np.int64).max() - 0) >= min(P, 1)
    np.testing.assert_array_equal(_np(state_array(state, "ranges")).view(np.uint32)[:G], o["ranges"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list")).view(np.uint32)[:K], o["point_list"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list_keys")).view(np.uint64)[:K], o["keys"])


def test_empty_and_all_culled():
    cam = scene.workload_camera("tiny")
    st, d, out = cuda_forward(cam, {k: v[:0] for k, v in scene.make_scene("tiny").items()}, bg=(0.2, 0.4, 0.6))
    color, depth, third, radii, _, state = out
    assert radii.numel() == 0 and state.num_rendered == 0
    assert torch.allclose(color[:, 0, 0].cpu(), torch.tensor([0.2, 0.4, 0.6]))
    assert float(depth.abs().max()) == 0 and float(third.abs().max()) == 0
    sc = scene.make_scene("tiny", P=100, behind=1.0, outside=0.0)      # everything behind the camera
    st, d, out = cuda_forward(cam, sc)
    assert int(out[3].abs().sum()) == 0 and out[5].num_rendered == 0
    got = rasterize_backward(st, out[5], d["means3D"], d["opacities"], d["shs"], None, d["scales"],
                             d["rotations"], None, torch.ones(3, cam.H, cam.W).cuda(),
                             torch.ones(1, cam.H, cam.W).cuda(), None)
    assert float(got["means3D"].abs().max()) == 0


def test_capacity_overflow_is_reissued(monkeypatch):
    """K above the binning capacity: render kernels exit, the binding re-issues stage 2."""
    from eags_raster import rasterize as rz
    sc = scene.make_scene("c1", P=40_000, seed=3, scale_range=(3.0, 6.0))
    cam = scene.workload_camera("c1")
    o = oracle_forward(cam, sc, skip_blend=True)
    assert o["num_rendered"] > (1 << 16)
    monkeypatch.setattr(rz, "_capacity", {torch.cuda.current_device(): 1 << 16})
    st, d, out = cuda_forward(cam, sc)
    assert out[5].num_rendered == o["num_rendered"] and out[5].capacity >= o["num_rendered"]
    np.testing.assert_array_equal(_np(state_array(out[5], "point_list")).view(np.uint32)[:o["num_rendered"]],
                                  o["point_list"])
    for mode in ("exact", "optimistic"):
        rz.set_sync_mode(mode)
        try:
            _, _, out2 = cuda_forward(cam, sc)
            assert torch.equal(out2[0], out[0]) and torch.equal(out2[1], out[1])
        finally:
            rz.set_sync_mode("optimistic")
