"""Pins of the tracking pose transform by the reference's own functions (tests/golden/tracking_reference.npz,
made by make_golden_tracking.py from src/utils/tracker_utils.py + gaussian_model_utils.build_rotation
combined as in src/entities/tracker.py:85-97).

CPU part: the "reference formulation" the GPU tests of the fast path compare against
(test_gpu_shims._quat_mul / _quat_to_R) is the reference's.  GPU part: eags_pose_compose folds the same
relative pose into the view matrices (viewmatrix' = Rel^T viewmatrix, i.e. w2c' = w2c Rel)."""
import os

import numpy as np
import pytest
import torch

from eags_raster import scene
from test_gpu_shims import _quat_mul, _quat_to_R, _settings_a

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tracking_reference.npz"))


def test_quaternion_product_order_is_the_reference_s():
    qa, qb = torch.from_numpy(G["qa"]), torch.from_numpy(G["qb"])
    # multiply_quaternions(q, r) is the Hamilton product r (x) q (tracker_utils.py:8-28)
    np.testing.assert_allclose(_quat_mul(qb, qa).numpy(), G["qa_times_qb_ref"], rtol=1e-6, atol=1e-6)
    assert np.abs(_quat_mul(qa, qb).numpy() - G["qa_times_qb_ref"]).max() > 1e-2


@pytest.mark.parametrize("c", range(6))
def test_test_side_pose_formulation_equals_tracker_py(c):
    pts, rots = torch.from_numpy(G["pts"]), torch.from_numpy(G["rots"])
    cr, ct = torch.from_numpy(G["cam_rot"][c]), torch.from_numpy(G["cam_trans"][c])
    R = _quat_to_R(cr)
    np.testing.assert_allclose(R.numpy(), G["rel_transform"][c][:3, :3], rtol=0, atol=2e-6)
    np.testing.assert_allclose((pts @ R.T + ct).numpy(), G["transformed_pts"][c], rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(_quat_mul((cr / cr.norm())[None], rots).numpy(), G["moved_rotations"][c],
                               rtol=1e-5, atol=2e-6)
    # compute_camera_opt_params returns the canonical (w >= 0) wxyz quaternion of the same rotation
    q = torch.from_numpy(G["opt_rot"][c])
    assert q[0] >= 0 and abs(float(q.norm()) - 1) < 1e-6
    np.testing.assert_allclose(_quat_to_R(q).numpy(), G["rel_transform"][c][:3, :3], rtol=0, atol=2e-6)
    np.testing.assert_allclose(G["opt_trans"][c], G["cam_trans"][c], rtol=0, atol=1e-7)
    np.testing.assert_allclose(G["rt_quat"][c], np.concatenate([G["opt_rot"][c], G["opt_trans"][c]]), atol=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize("c", range(6))
def test_pose_compose_kernel_against_reference_rel_transform(c):
    from eags_raster.tracking import compose_settings
    cam = scene.workload_camera("ragged", w2c=scene.rot_xy(3, -2))
    s = _settings_a(cam)
    rel = G["rel_transform"][c].astype(np.float64)
    s2, _ = compose_settings(s, torch.from_numpy(G["cam_rot"][c]).cuda(), torch.from_numpy(G["cam_trans"][c]).cuda())
    vm = cam.viewmatrix.numpy().astype(np.float64)      # = w2c^T (utils.py:97)
    pm = cam.projmatrix.numpy().astype(np.float64)      # = viewmatrix P^T (utils.py:98-104)
    want_vm = rel.T @ vm
    scale = max(1.0, np.abs(want_vm).max())
    np.testing.assert_allclose(s2.viewmatrix.cpu().numpy(), want_vm, rtol=0, atol=1e-5 * scale)
    want_pm = rel.T @ pm
    np.testing.assert_allclose(s2.projmatrix.cpu().numpy(), want_pm, rtol=0, atol=1e-5 * max(1.0, np.abs(want_pm).max()))
    want_campos = np.linalg.inv(want_vm.T)[:3, 3]
    np.testing.assert_allclose(s2.campos.cpu().numpy(), want_campos, rtol=0, atol=1e-5 * max(1.0, np.abs(want_campos).max()))
