"""Properties the domain gives for free, checked on the CPU oracle (the checker of every GPU parity test; the
reference's kernels are un-vendored, so besides the golden helper vectors and the float64 autograd restatement
these invariants are what ties the oracle to the rendering model rather than to its own code):

* the order in which the Gaussians are stored does not matter (binning sorts by tile and depth);
* a Gaussian with zero opacity, or behind the camera, contributes nothing;
* moving the camera and the scene by the same rigid transform leaves the image unchanged - with view-dependent
  colours (SH degree 3) too, which ties viewmatrix, projmatrix, campos and the quaternion convention together;
* the backward pass is linear in the upstream gradients;
* variant A's alpha and variant B's opacity are both 1 - final_T."""
import numpy as np
import torch

import oracle
from eags_raster import scene
from helpers import oracle_forward, rel_err
from test_gpu_shims import _quat_mul


def _scene(P=900, M=1, seed=4):
    return scene.make_scene("ragged", P=P, seed=seed, M=M)


def _ok(*os_):
    ok = np.ones_like(os_[0]["fragile"], bool)
    for o in os_:
        ok &= o["fragile"] == 0
    return ok


def test_storage_order_of_the_gaussians_does_not_matter():
    sc = _scene()
    cam = scene.workload_camera("ragged", w2c=scene.rot_xy(2, -1))
    a = oracle_forward(cam, sc, bg=(0.1, 0.2, 0.3))
    perm = torch.randperm(sc["means3D"].shape[0], generator=torch.Generator().manual_seed(1))
    sp = {k: (v[perm].contiguous() if v.shape[0] == perm.numel() and k not in ("gt_color", "gt_depth") else v)
          for k, v in sc.items()}
    b = oracle_forward(cam, sp, bg=(0.1, 0.2, 0.3))
    assert np.array_equal(a["radii"][perm.numpy()], b["radii"])
    assert a["num_rendered"] == b["num_rendered"] and np.array_equal(a["ranges"], b["ranges"])
    # same depth order per tile (no exact depth ties in this scene): the sorted lists name the same Gaussians
    assert np.array_equal(perm.numpy()[b["point_list"]], a["point_list"])
    assert np.array_equal(a["n_contrib"], b["n_contrib"])
    for k in ("color", "depth", "third", "final_T"):
        assert np.array_equal(a[k], b[k]), k            # identical summation order -> identical bits


def test_transparent_and_behind_camera_gaussians_contribute_nothing():
    sc = _scene()
    cam = scene.workload_camera("ragged")
    P = sc["means3D"].shape[0]
    kill = torch.zeros(P, dtype=torch.bool)
    kill[::3] = True
    a = dict(sc)
    a["opacities"] = torch.where(kill[:, None], torch.zeros_like(sc["opacities"]), sc["opacities"])
    b = {k: (v[~kill].contiguous() if v.shape[0] == P and k not in ("gt_color", "gt_depth") else v) for k, v in sc.items()}
    oa, ob = oracle_forward(cam, a), oracle_forward(cam, b)
    ok = _ok(oa, ob)
    for k in ("color", "depth", "third"):
        assert np.abs(oa[k][:, ok] - ob[k][:, ok]).max() == 0, k
    c = dict(sc)
    c["means3D"] = sc["means3D"].clone()
    c["means3D"][kill, 2] = -c["means3D"][kill, 2].abs()          # behind the camera: culled (radius 0)
    oc = oracle_forward(cam, c)
    assert (oc["radii"][kill.numpy()] == 0).all()
    ok = _ok(oc, ob)
    for k in ("color", "depth", "third"):
        assert np.abs(oc[k][:, ok] - ob[k][:, ok]).max() == 0, k


def test_rigid_motion_of_camera_and_scene_together_leaves_the_image():
    sc = _scene(P=700, M=16, seed=6)
    cam0 = scene.workload_camera("ragged", w2c=scene.rot_xy(1, 2))
    # world motion X' = R X + t;  camera: w2c' = w2c @ inv([R t])
    T = scene.rot_xy(20, -35).double()
    T[:3, 3] = torch.tensor([0.4, -0.3, 0.7], dtype=torch.float64)
    R = T[:3, :3]
    from scipy.spatial.transform import Rotation
    q_xyzw = Rotation.from_matrix(R.numpy()).as_quat()
    qR = torch.tensor([q_xyzw[3], q_xyzw[0], q_xyzw[1], q_xyzw[2]], dtype=torch.float32)
    moved = dict(sc)
    moved["means3D"] = (sc["means3D"].double() @ R.T + T[:3, 3]).float()
    moved["rotations"] = _quat_mul(qR[None], sc["rotations"])       # world rotation applied on the left
    cam1 = scene.workload_camera("ragged", w2c=(cam0.w2c.double() @ torch.linalg.inv(T)).float())
    # degree 0: colours do not depend on the view direction
    a, b = oracle_forward(cam0, sc, degree=0), oracle_forward(cam1, moved, degree=0)
    same = a["radii"] == b["radii"]
    assert same.mean() > 0.99                                       # a ceil() may flip under fp32 rounding
    ok = _ok(a, b)
    for k, tol in (("color", 2e-4), ("depth", 5e-4), ("third", 2e-4)):
        err = np.abs(a[k][:, ok] - b[k][:, ok])
        assert np.median(err) < 1e-6 and (err > tol).mean() < 5e-3, (k, float(err.max()))
    # degree 3 with the SH frame fixed to the world: rotating the world rotates the view directions, so only a pure
    # translation leaves view-dependent colours unchanged
    Tt = torch.eye(4, dtype=torch.float64)
    Tt[:3, 3] = T[:3, 3]
    shifted = dict(sc)
    shifted["means3D"] = (sc["means3D"].double() + Tt[:3, 3]).float()
    cam2 = scene.workload_camera("ragged", w2c=(cam0.w2c.double() @ torch.linalg.inv(Tt)).float())
    a3, b3 = oracle_forward(cam0, sc, degree=3), oracle_forward(cam2, shifted, degree=3)
    ok = _ok(a3, b3)
    err = np.abs(a3["color"][:, ok] - b3["color"][:, ok])
    assert np.median(err) < 1e-6 and (err > 2e-4).mean() < 5e-3, float(err.max())
    assert np.abs(a3["color"] - a["color"]).max() > 1e-3            # the higher SH bands do change the colours


def test_backward_is_linear_in_the_upstream_gradients():
    sc = _scene(P=600, M=4, seed=8)
    cam = scene.workload_camera("ragged", w2c=scene.rot_xy(-2, 3))
    o = oracle_forward(cam, sc, degree=1, variant=1)
    g = torch.Generator().manual_seed(2)
    ups = [[torch.randn(c, cam.H, cam.W, generator=g).numpy() for c in (3, 1, 1)] for _ in range(2)]
    ga = oracle.backward(o, *ups[0], want_tau=True)
    gb = oracle.backward(o, *ups[1], want_tau=True)
    mix = [2.0 * x - 0.5 * y for x, y in zip(*ups)]
    gm = oracle.backward(o, *mix, want_tau=True)
    for k in ("dL_dmeans3D", "dL_dscales", "dL_drots", "dL_dopacity", "dL_dshs", "dL_dmean2D", "dL_dtau"):
        assert rel_err(gm[k], 2.0 * ga[k].astype(np.float64) - 0.5 * gb[k].astype(np.float64)) < 2e-5, k


def test_third_output_is_one_minus_final_T_in_both_variants():
    sc = _scene(P=500, seed=9)
    cam = scene.workload_camera("ragged")
    for variant in (0, 1):
        o = oracle_forward(cam, sc, variant=variant)
        np.testing.assert_allclose(o["third"][0], 1.0 - o["final_T"], rtol=0, atol=2e-6)
    a, b = oracle_forward(cam, sc, variant=0), oracle_forward(cam, sc, variant=1)
    assert np.array_equal(a["color"], b["color"]) and np.array_equal(a["radii"], b["radii"])
