"""Per-view settings against vectors made by the reference's own get_render_settings
(src/utils/utils.py:77-117) importing THIS repo's gaussian_rasterizer shim as its rasterizer package
(tests/golden/make_golden_settings.py).  scene.make_camera - the camera every parity test, the oracle and
bench.py use - must build the same view / projection matrices, camera centre and tan(fov)."""
import os

import numpy as np
import pytest
import torch

from eags_raster import scene

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "settings_reference.npz"))


@pytest.mark.parametrize("i", range(4))
def test_make_camera_equals_get_render_settings(i):
    W, H, fx, fy, cx, cy = G["cams"][i]
    W, H = int(W), int(H)
    cam = scene.make_camera(W, H, fx, fy, cx, cy, w2c=torch.from_numpy(G[f"{i}/w2c"]))
    sc = G[f"{i}/scalars"]
    assert (cam.H, cam.W) == (int(sc[0]), int(sc[1]))
    assert abs(cam.tanfovx - sc[2]) <= 1e-12 and abs(cam.tanfovy - sc[3]) <= 1e-12
    assert sc[4] == 1.0 and sc[5] == 0 and sc[6] == 0 and sc[7] == 0          # scale_modifier, sh_degree, flags
    np.testing.assert_array_equal(cam.viewmatrix.numpy(), G[f"{i}/viewmatrix"])
    np.testing.assert_allclose(cam.projmatrix.numpy(), G[f"{i}/projmatrix"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(cam.campos.numpy(), G[f"{i}/campos"], rtol=1e-6, atol=1e-7)
    np.testing.assert_array_equal(G[f"{i}/bg"], np.zeros(3, np.float32))


def test_shim_settings_take_the_reference_s_keywords():
    from gaussian_rasterizer import GaussianRasterizationSettings
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")          # utils.py:105-117, in order


@pytest.mark.parametrize("i", range(4))
def test_variant_b_camera_equals_gsr_camera(i):
    """src/gsr/camera.py:97-111 + lc.py:70-77 (projection built with fy := fx, fov from the real fy): the same
    w2c-transposed / view @ proj_raw / inverse-row conventions as variant A, so scene.make_camera(fy=fx) gives
    the matrices and the true fy gives tanfovy."""
    W, H, fx, fy, cx, cy = G["cams"][i]
    W, H = int(W), int(H)
    w2c = torch.from_numpy(G[f"{i}/w2c"])
    quirk = scene.make_camera(W, H, fx, fx, cx, cy, w2c=w2c)        # matrices: fy := fx
    true = scene.make_camera(W, H, fx, fy, cx, cy, w2c=w2c)         # tan(fov): real fy
    np.testing.assert_allclose(G[f"{i}/b_tanfov"], [true.tanfovx, true.tanfovy], rtol=1e-12)
    np.testing.assert_allclose(quirk.viewmatrix.numpy(), G[f"{i}/b_viewmatrix"], rtol=0, atol=1e-7)
    np.testing.assert_allclose(quirk.projmatrix_raw.numpy(), G[f"{i}/b_projmatrix_raw"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(quirk.projmatrix.numpy(), G[f"{i}/b_projmatrix"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(quirk.campos.numpy(), G[f"{i}/b_campos"], rtol=1e-6, atol=1e-6)
