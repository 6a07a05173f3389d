"""Full-size checks of the CUDA path (BASELINE.json shapes c2-c4) through size-independent
properties, plus one full-size parity run against the oracle at the headline shape c2.

Properties (all through the C ABI binding):
  * binning: sum(tiles_touched) == K, ranges partition [0, K) in tile order, every tile's keys carry
    its tile id and are sorted by depth bits with ties in Gaussian-index order (== the stable sort),
    radii > 0 <=> tiles_touched > 0;
  * forward determinism (bit-identical reruns), background linearity colour(bg) = colour(0) + T bg,
    variant A alpha == variant B opacity == 1 - T where the pixel did not terminate early,
    dropping the culled Gaussians changes nothing (bit-identical images);
  * backward linearity in dL/dpixels, exact directional derivative w.r.t. SH colours (the forward is
    linear in them), zero gradients for culled Gaussians.
"""
import numpy as np
import pytest
import torch

import oracle
from eags_raster import rasterize_backward, scene, state_array
from helpers import assert_close_rel, cuda_forward, nonfragile, oracle_forward, rel_err

pytestmark = pytest.mark.gpu


def _np(t):
    return t.detach().cpu().numpy()


SHAPES = [("c2", None), ("c3", None), ("c4", None)]


@pytest.mark.parametrize("wl,P", SHAPES)
def test_binning_invariants_full_size(wl, P):
    sc = scene.make_scene(wl, P=P, seed=0)
    cam = scene.workload_camera(wl, w2c=scene.rot_xy(2, -3))
    st, d, out = cuda_forward(cam, sc)
    color, depth, third, radii, _, state = out
    K = state.num_rendered
    tt = state_array(state, "tiles_touched").to(torch.int64)
    assert int(tt.sum()) == K and K > 2 * state.P
    assert torch.equal(radii > 0, tt > 0)
    gx, gy = (cam.W + 15) // 16, (cam.H + 15) // 16
    G = gx * gy
    rg = state_array(state, "ranges").to(torch.int64)[:G]
    cnt = rg[:, 1] - rg[:, 0]
    assert int(cnt.sum()) == K and int(cnt.min()) >= 0
    nz = cnt > 0
    starts = torch.cumsum(cnt, 0) - cnt
    assert torch.equal(rg[nz, 0], starts[nz]) and int(rg[~nz].abs().sum()) == 0
    keys = state_array(state, "point_list_keys")[:K]
    vals = state_array(state, "point_list")[:K].to(torch.int64)
    tile_of_key = keys >> 32
    # keys sorted globally as unsigned 64-bit (top bit is clear: tile ids < 2^31)
    assert bool((keys[1:] >= keys[:-1]).all())
    assert torch.equal(torch.bincount(tile_of_key, minlength=G)[:G], cnt)
    # ties (same tile, same depth bits) keep Gaussian-index order == emission order of a stable sort
    tie = keys[1:] == keys[:-1]
    assert bool((vals[1:][tie] > vals[:-1][tie]).all())
    # the key's depth bits are the Gaussian's view depth
    depth_bits = state_array(state, "rgbd")[:, 3].contiguous().view(torch.int32).to(torch.int64)
    assert torch.equal(keys & 0xFFFFFFFF, depth_bits[vals])
    # every (tile, Gaussian) pair lies in the Gaussian's tile rectangle count
    assert torch.equal(torch.bincount(vals, minlength=state.P), tt)
    assert torch.isfinite(color).all() and torch.isfinite(depth).all() and torch.isfinite(third).all()
    nc = state_array(state, "n_contrib").to(torch.int64)
    per_tile = cnt.view(gy, gx).repeat_interleave(16, 0).repeat_interleave(16, 1)[: cam.H, : cam.W]
    assert bool((nc <= per_tile).all())


@pytest.mark.parametrize("wl", ["c2", "c3"])
def test_forward_properties_full_size(wl):
    sc = scene.make_scene(wl, seed=1)
    cam = scene.workload_camera(wl)
    bg = (0.25, 0.5, 0.75)
    _, _, a0 = cuda_forward(cam, sc, with_keys=False)
    _, _, a1 = cuda_forward(cam, sc, with_keys=False)
    for i in range(4):                                  # determinism: bit-identical reruns
        assert torch.equal(a0[i], a1[i])
    _, _, ab = cuda_forward(cam, sc, bg=bg, with_keys=False)
    T = state_array(a0[5], "final_T")
    want = a0[0] + T[None] * torch.tensor(bg, device=T.device)[:, None, None]
    assert float((ab[0] - want).abs().max()) <= 1e-6
    assert torch.equal(ab[1], a0[1])
    # variant B: same colour / depth; opacity = 1 - T; equals alpha where blending did not stop early
    _, _, b0 = cuda_forward(cam, sc, variant=1, with_keys=False)
    assert torch.equal(b0[0], a0[0]) and torch.equal(b0[1], a0[1])
    assert float((b0[2][0] - (1 - T)).abs().max()) <= 1e-6
    assert float((a0[2][0] - (1 - T)).abs().max()) <= 2e-5
    nt = b0[4]
    assert int(nt.min()) >= 0 and int(nt[a0[3] == 0].sum()) == 0 and int(nt.sum()) > 0
    # dropping culled Gaussians changes nothing
    keep = (a0[3] > 0).cpu()
    sub = {k: (v[keep] if v.shape[0] == keep.shape[0] else v) for k, v in sc.items()}
    _, _, s0 = cuda_forward(cam, sub, with_keys=False)
    assert torch.equal(s0[0], a0[0]) and torch.equal(s0[1], a0[1]) and torch.equal(s0[2], a0[2])


@pytest.mark.parametrize("wl", ["c2", "c4"])
def test_backward_properties_full_size(wl):
    sc = scene.make_scene(wl, seed=2)
    cam = scene.workload_camera(wl)
    st, d, out = cuda_forward(cam, sc, with_keys=False)
    state = out[5]
    H, W = cam.H, cam.W
    g = torch.Generator(device="cuda").manual_seed(3)
    G1 = [torch.randn(3, H, W, device="cuda", generator=g), torch.randn(1, H, W, device="cuda", generator=g),
          torch.randn(1, H, W, device="cuda", generator=g)]
    G2 = [torch.randn(3, H, W, device="cuda", generator=g), torch.randn(1, H, W, device="cuda", generator=g),
          torch.randn(1, H, W, device="cuda", generator=g)]

    def bwd(gs):
        return rasterize_backward(st, state, d["means3D"], d["opacities"], d["shs"], None, d["scales"],
                                  d["rotations"], None, gs[0], gs[1], gs[2], want_tau=True)

    r1, r2 = bwd(G1), bwd(G2)
    r12 = bwd([0.5 * a - 2.0 * b for a, b in zip(G1, G2)])
    for k in ("means3D", "means2D", "opacities", "shs", "scales", "rotations", "tau"):
        want = 0.5 * r1[k] - 2.0 * r2[k]
        assert rel_err(_np(r12[k]), _np(want)) < 1e-3, k            # linear in dL/dpixels (fp32 atomics order)
        assert torch.isfinite(r12[k]).all(), k
    cul = out[3] == 0
    assert float(r1["means3D"][cul].abs().max()) == 0 and float(r1["shs"][cul].abs().max()) == 0
    # colour is linear in the SH-DC coefficients (where not clamped): <dL/dshs, dir> == L(shs + dir) - L(shs)
    direction = 0.05 * torch.randn(d["shs"].shape, device="cuda", generator=g)
    # stay on one side of the max(0, .) clamp: perturb only channels at least 0.1 above it
    direction = torch.where(0.28209479177387814 * d["shs"] + 0.5 > 0.1, direction, torch.zeros_like(direction))
    sc2 = dict(sc)
    sc2["shs"] = (d["shs"] + direction).cpu()
    _, _, out2 = cuda_forward(cam, sc2, with_keys=False)
    vis = out[3] > 0      # the clamp flags are only written for visible Gaussians
    assert torch.equal(state_array(state, "clamped")[: state.P][vis], state_array(out2[5], "clamped")[: state.P][vis])
    lhs = float(((out2[0] - out[0]).double() * G1[0].double()).sum())
    rhs = float((r1["shs"].double() * direction.double()).sum())
    scale = float(((out2[0] - out[0]).double() * G1[0].double()).abs().sum())
    assert abs(lhs - rhs) <= 1e-4 * scale + 1e-6, (lhs, rhs, scale)


def test_full_size_parity_with_oracle_c2():
    """The headline shape itself against the oracle (the C oracle needs ~1 s for it)."""
    sc = scene.make_scene("c2", seed=0)
    cam = scene.workload_camera("c2")
    o = oracle_forward(cam, sc)
    st, d, out = cuda_forward(cam, sc)
    color, depth, third, radii, _, state = out
    K = o["num_rendered"]
    np.testing.assert_array_equal(_np(radii), o["radii"])
    assert state.num_rendered == K
    np.testing.assert_array_equal(_np(state_array(state, "ranges")).view(np.uint32)[: o["ranges"].shape[0]], o["ranges"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list")).view(np.uint32)[:K], o["point_list"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list_keys")).view(np.uint64)[:K], o["keys"])
    ok = nonfragile(o)
    assert ok.mean() >= 0.999, f"fragile fraction {1 - ok.mean():.4%}"
    np.testing.assert_array_equal(_np(state_array(state, "n_contrib")).view(np.uint32)[ok], o["n_contrib"][ok])
    assert_close_rel(_np(color)[:, ok], o["color"][:, ok], 1e-5, 1e-5, "color")
    assert_close_rel(_np(depth)[0][ok], o["depth"][0][ok], 1e-5, 1e-5, "depth")
    assert_close_rel(_np(third)[0][ok], o["third"][0][ok], 1e-5, 1e-5, "alpha")
    H, W = cam.H, cam.W
    dC = np.sign(o["color"] - sc["gt_color"].numpy()).astype(np.float32) / (3 * H * W)
    dD = np.sign(o["depth"] - sc["gt_depth"].numpy()).astype(np.float32) / (H * W)
    dC[:, ~ok] = 0
    dD[:, ~ok] = 0
    ref = oracle.backward(o, dC, dD, None, want_tau=True)
    got = rasterize_backward(st, state, d["means3D"], d["opacities"], d["shs"], None, d["scales"],
                             d["rotations"], None, torch.from_numpy(dC).cuda(), torch.from_numpy(dD).cuda(),
                             None, want_tau=True)
    assert rel_err(_np(got["means3D"]), ref["dL_dmeans3D"]) < 1e-3
    assert rel_err(_np(got["means2D"])[:, :2], ref["dL_dmean2D"]) < 1e-3
    assert rel_err(_np(got["scales"]), ref["dL_dscales"]) < 1e-3
    assert rel_err(_np(got["rotations"]), ref["dL_drots"]) < 1e-3
    assert rel_err(_np(got["opacities"]).reshape(-1), ref["dL_dopacity"]) < 1e-3
    assert rel_err(_np(got["shs"]), ref["dL_dshs"]) < 1e-3
    assert rel_err(_np(got["tau"]), ref["dL_dtau"]) < 1e-3
