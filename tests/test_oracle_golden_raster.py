"""The CPU oracle reproduces its own frozen vectors (tests/golden/raster_oracle.npz): integer / index /
key outputs bit for bit, floats to 1e-6 relative (the oracle is compiled without FMA contraction, so a
change here means the checker's arithmetic moved - compiler, flags or an edit of eags_oracle.c)."""
import numpy as np
import pytest

import oracle
from golden_raster import CASES, load
from helpers import oracle_forward

INT_KEYS = ("radii", "tiles_touched", "ranges", "point_list", "keys", "clamped", "n_contrib", "fragile")
F_KEYS = ("xy", "depths", "rgb", "conic_opacity", "color", "depth", "third", "final_T")
G_KEYS = ("dL_dmean2D", "dL_dmeans3D", "dL_dscales", "dL_drots", "dL_dopacity", "dL_dshs", "dL_dtau")


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_frozen_vectors(name):
    g, cam, sc, meta = load(name)
    o = oracle_forward(cam, sc, degree=meta["degree"], bg=meta["bg"], variant=meta["variant"])
    assert o["num_rendered"] == meta["K"]
    for k in INT_KEYS:
        np.testing.assert_array_equal(o[k], g[k], err_msg=k)
    for k in F_KEYS:
        np.testing.assert_allclose(o[k], g[k], rtol=1e-6, atol=1e-7, err_msg=k)
    if meta["variant"] == 1:
        np.testing.assert_array_equal(o["n_touched"], g["n_touched"])
    b = oracle.backward(o, g["dL_dcolor"], g["dL_ddepth"], g["dL_dthird"], want_tau=True)
    for k in G_KEYS:
        scale = max(float(np.abs(g[k]).max()), 1e-30)
        assert float(np.abs(b[k] - g[k]).max()) <= 1e-5 * scale, k


def test_frozen_vectors_are_non_trivial():
    for name in CASES:
        g, cam, sc, meta = load(name)
        assert meta["K"] > 1000 and (g["radii"] > 0).sum() > 100 and (g["radii"] == 0).sum() > 0
        assert (g["fragile"] == 0).mean() > 0.98
        assert g["ranges"].shape[0] == ((cam.W + 15) // 16) * ((cam.H + 15) // 16)
        assert np.abs(g["dL_dtau"]).max() > 0
