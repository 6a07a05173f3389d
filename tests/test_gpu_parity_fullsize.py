"""Full-size oracle parity on every BASELINE.json config and on the shapes the reference really calls with.

VERDICT r1 item 2: c2 (both variants), c3 and c4 at full size, plus SURVEY Appendix B's call shapes
(TUM 380x540 and ScanNet 456x616 after the crop, ScanNet++ 584x876, Replica 680x1200) and the
M = 16 / degree 0 stride of evaluate_merged_map.py:67.  Bars (BASELINE.json north_star), written here:

  * radii, K, tile ranges, point_list, sorted 64-bit keys, n_contrib: bit-exact;
  * colour / depth / alpha / final_T: 1e-5 TRUE elementwise relative error wherever |ref| > 1e-3 (no
    slack borrowed from the brightest pixel), 1e-8 absolute below;
  * gradients (means3D, means2D, scales, rotations, opacities, shs, dL/dtau): 1e-3, checked ROW-wise --
    every Gaussian whose reference gradient exceeds 1e-3 of the largest is held to 1e-3 of its own norm.

Fragile pixels (a hard threshold of SURVEY A.6 Q6/Q10 within 1e-5 of flipping in the oracle's own
arithmetic; excluded from the float comparisons, their upstream gradients zeroed) must stay below 0.1 %
of the image; the fraction is printed.  They differ by one blended splat: the 1/255 alpha cut or the
T < 1e-4 stop lands on the other side in the other implementation's rounding.
"""
import numpy as np
import pytest
import torch

import oracle
from eags_raster import rasterize_backward, scene, state_array
from helpers import assert_pixels_rel, assert_rows_rel, cuda_forward, nonfragile, oracle_forward, rel_err

pytestmark = pytest.mark.gpu

RTOL_IMG = 1e-5
RTOL_GRAD = 1e-3

CASES = [
    # workload, P (None = the config's), M, variant, camera rotation (deg x, deg y)
    ("c2", None, 1, 0, (0, 0)),
    ("c2", None, 1, 1, (1, -1)),
    ("c3", None, 1, 0, (0, 0)),
    ("c4", None, 1, 0, (0, 0)),
    ("tum_crop", None, 1, 0, (2, 1)),
    ("scannet_crop", None, 1, 0, (-1, 2)),
    ("scannetpp", None, 1, 0, (0, -3)),
    ("replica", None, 16, 0, (1, 1)),      # M = 16 rows at degree 0: the stride of evaluate_merged_map.py:67
    ("tum_crop", 100_000, 16, 1, (0, 0)),
]


def _np(t):
    return t.detach().cpu().numpy()


@pytest.mark.parametrize("wl,P,M,variant,rot", CASES)
def test_full_size_parity(wl, P, M, variant, rot):
    sc = scene.make_scene(wl, P=P, seed=0, M=M)
    cam = scene.workload_camera(wl, w2c=scene.rot_xy(*rot))
    o = oracle_forward(cam, sc, variant=variant)
    st, d, out = cuda_forward(cam, sc, variant=variant)
    color, depth, third, radii, n_touched, state = out
    K, G = o["num_rendered"], o["grid"][0] * o["grid"][1]
    # ---- integers / keys: bit-exact
    np.testing.assert_array_equal(_np(radii), o["radii"])
    assert state.num_rendered == K
    np.testing.assert_array_equal(_np(state_array(state, "tiles_touched")).view(np.uint32)[: o["P"]], o["tiles_touched"])
    np.testing.assert_array_equal(_np(state_array(state, "ranges")).view(np.uint32)[:G], o["ranges"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list")).view(np.uint32)[:K], o["point_list"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list_keys")).view(np.uint64)[:K], o["keys"])
    ok = nonfragile(o)
    frag = 1.0 - float(ok.mean())
    assert frag < 1e-3, f"fragile pixels: {frag:.5%}"
    np.testing.assert_array_equal(_np(state_array(state, "n_contrib")).view(np.uint32)[ok], o["n_contrib"][ok])
    # ---- images: true elementwise relative error
    worst = {
        "color": assert_pixels_rel(_np(color)[:, ok], o["color"][:, ok], RTOL_IMG, 1e-3, "color"),
        "depth": assert_pixels_rel(_np(depth)[0][ok], o["depth"][0][ok], RTOL_IMG, 1e-3, "depth"),
        # variant A: alpha = sum alpha_i T_i (a sum of positive terms).  Variant B: opacity = 1 - T_final, a
        # difference of two O(1) numbers -- one ulp of T = 0.999 is already 6e-5 of an opacity of 1e-3 -- so
        # it is held to 1e-5 of the larger operand (1), and T itself to 1e-5 relative on the line below
        "third": assert_pixels_rel(_np(third)[0][ok], o["third"][0][ok], RTOL_IMG, 1e-3 if variant == 0 else 1.0,
                                   "alpha/opacity"),
        "final_T": assert_pixels_rel(_np(state_array(state, "final_T"))[ok], o["final_T"][ok], RTOL_IMG, 1e-3, "final_T"),
    }
    # ---- gradients of the L1 colour + depth loss of bench.py (+ a dense dL/dalpha for variant A)
    H, W = cam.H, cam.W
    dC = np.sign(o["color"] - sc["gt_color"].numpy()).astype(np.float32) / (3 * H * W)
    dD = np.sign(o["depth"] - sc["gt_depth"].numpy()).astype(np.float32) / (H * W)
    dA = None
    if variant == 0:
        dA = (np.random.default_rng(3).standard_normal((1, H, W)).astype(np.float32)) / (H * W)
        dA[:, ~ok] = 0
    dC[:, ~ok] = 0
    dD[:, ~ok] = 0
    ref = oracle.backward(o, dC, dD, dA, want_tau=True)
    got = rasterize_backward(st, state, d["means3D"], d["opacities"], d["shs"], None, d["scales"],
                             d["rotations"], None, torch.from_numpy(dC).cuda(), torch.from_numpy(dD).cuda(),
                             None if dA is None else torch.from_numpy(dA).cuda(), want_tau=True)
    rows = {
        "means3D": assert_rows_rel(_np(got["means3D"]), ref["dL_dmeans3D"], RTOL_GRAD, 1e-3, "dL/dmeans3D"),
        "means2D": assert_rows_rel(_np(got["means2D"])[:, :2], ref["dL_dmean2D"], RTOL_GRAD, 1e-3, "dL/dmeans2D"),
        "scales": assert_rows_rel(_np(got["scales"]), ref["dL_dscales"], RTOL_GRAD, 1e-3, "dL/dscales"),
        "rotations": assert_rows_rel(_np(got["rotations"]), ref["dL_drots"], RTOL_GRAD, 1e-3, "dL/drotations"),
        "opacities": assert_rows_rel(_np(got["opacities"]).reshape(-1, 1), ref["dL_dopacity"].reshape(-1, 1),
                                     RTOL_GRAD, 1e-3, "dL/dopacities"),
        "shs": assert_rows_rel(_np(got["shs"]), ref["dL_dshs"], RTOL_GRAD, 1e-3, "dL/dshs"),
    }
    assert rel_err(_np(got["tau"]), ref["dL_dtau"]) < RTOL_GRAD
    cul = o["radii"] == 0
    assert np.all(_np(got["means3D"])[cul] == 0) and np.all(_np(got["shs"])[cul] == 0)
    print(f"[{wl} {W}x{H} P={o['P']} M={M} variant={'AB'[variant]}] K={K} fragile {frag:.4%} | worst pixel rel "
          + " ".join(f"{k} {v:.1e}" for k, v in worst.items()) + " | worst row rel "
          + " ".join(f"{k} {v:.1e}" for k, v in rows.items())
          + f" | tau {rel_err(_np(got['tau']), ref['dL_dtau']):.1e}")
