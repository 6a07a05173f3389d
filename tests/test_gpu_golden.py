"""The CUDA path (through the C ABI) against the committed vectors of tests/golden/raster_oracle.npz -
no oracle call at run time.  Same bars as test_gpu_parity.py: radii, tiles, ranges, point list, sorted
keys, pixel means, depth keys and n_contrib bit-exact; images 1e-5 relative; gradients 1e-3 relative."""
import numpy as np
import pytest
import torch

from eags_raster import rasterize_backward, state_array
from golden_raster import CASES, load
from helpers import assert_close_rel, cuda_forward, rel_err

pytestmark = pytest.mark.gpu

RTOL_IMG = 1e-5
RTOL_GRAD = 1e-3


def _np(t):
    return t.detach().cpu().numpy()


@pytest.mark.parametrize("name", CASES)
def test_cuda_matches_frozen_vectors(name):
    g, cam, sc, meta = load(name)
    st, d, out = cuda_forward(cam, sc, degree=meta["degree"], bg=meta["bg"], variant=meta["variant"])
    color, depth, third, radii, n_touched, state = out
    P, K = sc["means3D"].shape[0], meta["K"]
    G = g["ranges"].shape[0]
    np.testing.assert_array_equal(_np(radii), g["radii"])
    assert state.num_rendered == K
    np.testing.assert_array_equal(_np(state_array(state, "tiles_touched")).view(np.uint32)[:P], g["tiles_touched"])
    np.testing.assert_array_equal(_np(state_array(state, "ranges")).view(np.uint32)[:G], g["ranges"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list")).view(np.uint32)[:K], g["point_list"])
    np.testing.assert_array_equal(_np(state_array(state, "point_list_keys")).view(np.uint64)[:K], g["keys"])
    vis = g["radii"] > 0
    np.testing.assert_array_equal(_np(state_array(state, "xy"))[:P][vis], g["xy"][vis])
    np.testing.assert_array_equal(_np(state_array(state, "rgbd"))[:P][vis, 3], g["depths"][vis])
    ok = g["fragile"] == 0
    np.testing.assert_array_equal(_np(state_array(state, "n_contrib")).view(np.uint32)[ok], g["n_contrib"][ok])
    assert_close_rel(_np(color)[:, ok], g["color"][:, ok], 1e-5, RTOL_IMG, "color")
    assert_close_rel(_np(depth)[0][ok], g["depth"][0][ok], 1e-5, RTOL_IMG, "depth")
    assert_close_rel(_np(third)[0][ok], g["third"][0][ok], 1e-5, RTOL_IMG, "alpha/opacity")
    if meta["variant"] == 1:
        okg = g["n_touched_fragile"] == 0
        assert ((_np(n_touched) != g["n_touched"]) & okg).sum() <= 4 * (~ok).sum()
    dC, dD, dA = (torch.from_numpy(g[k].copy()).cuda() for k in ("dL_dcolor", "dL_ddepth", "dL_dthird"))
    got = rasterize_backward(st, state, d["means3D"], d["opacities"], d["shs"], None, d["scales"], d["rotations"],
                             None, dC, dD, dA, want_tau=True)
    assert rel_err(_np(got["means2D"])[:, :2], g["dL_dmean2D"]) < RTOL_GRAD
    assert rel_err(_np(got["means3D"]), g["dL_dmeans3D"]) < RTOL_GRAD
    assert rel_err(_np(got["scales"]), g["dL_dscales"]) < RTOL_GRAD
    assert rel_err(_np(got["rotations"]), g["dL_drots"]) < RTOL_GRAD
    assert rel_err(_np(got["opacities"]).reshape(-1), g["dL_dopacity"]) < RTOL_GRAD
    assert rel_err(_np(got["shs"]), g["dL_dshs"]) < RTOL_GRAD
    assert rel_err(_np(got["tau"]), g["dL_dtau"]) < RTOL_GRAD
