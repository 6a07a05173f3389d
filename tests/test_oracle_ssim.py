"""CPU: the SSIM oracle (oracle/ssim.py) against golden vectors produced by the reference's own
``ssim`` and its autograd gradient (tests/golden/ssim_reference.npz, made by make_golden_ssim.py from
/root/reference/src/entities/losses.py).  The reference computes in fp32: tolerance 2e-6 absolute on the
value, 1e-4 of the gradient's maximum."""
import os

import numpy as np
import pytest

from oracle import ssim as oss

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ssim_reference.npz"))


@pytest.mark.parametrize("name", ["ragged", "square", "gray"])
def test_oracle_matches_reference_ssim(name):
    val, grad = oss.ssim_and_grad(G[f"{name}_img"], G[f"{name}_gt"])
    assert abs(val - float(G[f"{name}_ssim"])) < 2e-6
    ref = G[f"{name}_grad"]
    assert np.abs(grad - ref).max() < 1e-4 * np.abs(ref).max()


def test_window_is_the_references_float32_gaussian():
    g = oss.window()
    assert g.shape == (11,) and abs(g.sum() - 1) < 1e-6 and np.allclose(g, g[::-1]) and g.argmax() == 5


def test_isotropic_golden_is_self_consistent():
    """the golden isotropic loss equals its definition in float64 (guards the fixture itself)"""
    R = np.load(os.path.join(os.path.dirname(__file__), "golden", "isotropic_reference.npz"))
    s = R["scaling"].astype(np.float64)
    d = s - s.mean(axis=1, keepdims=True)
    assert abs(np.abs(d).mean() - float(R["loss"])) < 1e-6 * float(R["loss"])
    sg = np.sign(d)
    want = 2.0 * (sg - sg.mean(axis=1, keepdims=True)) / s.size
    big = np.abs(d) > 1e-9           # rows that are exactly isotropic sit on the kink of |.|
    assert np.abs(want - R["grad"])[big].max() < 1e-6 * np.abs(want).max()
