"""Tracking fast path (eags_raster.tracking.render_relative) against the
reference's formulation: move all Gaussians by the relative pose in PyTorch (tracker.py:85-97),
render them with the ordinary variant-A operator and back-propagate through dL/dmeans3D and
dL/drotations.  Images agree to fp32 rounding, the 7 pose gradients within 1e-3 relative."""
import numpy as np
import pytest
import torch

from eags_raster import scene
from test_gpu_shims import _quat_mul, _quat_to_R, _settings_a

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("wl,P,rot", [("ragged", 3000, (2, -1)), ("c1", 50_000, (0, 0))])
def test_render_relative_matches_transformed_gaussians(wl, P, rot):
    from eags_raster.losses import tracking_losses
    from eags_raster.tracking import render_relative
    from gaussian_rasterizer import GaussianRasterizer
    sc = scene.make_scene(wl, P=P, seed=5)
    cam = scene.workload_camera(wl, w2c=scene.rot_xy(*rot))
    s = _settings_a(cam)
    base = {k: sc[k].to(DEV) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    gt_c, gt_d = sc["gt_color"].to(DEV), sc["gt_depth"].to(DEV)
    mask = torch.ones(1, cam.H, cam.W, dtype=torch.bool, device=DEV)

    def loss_of(color, depth, alpha):
        lc, ld = tracking_losses(color, depth, alpha, gt_c, gt_d, mask, soft_alpha=True)
        return 0.6 * lc + 0.4 * ld

    q0 = [1.0, 0.006, -0.004, 0.003]
    t0 = [0.012, -0.007, 0.009]
    # reference formulation
    cr = torch.tensor(q0, device=DEV, requires_grad=True)
    ct = torch.tensor(t0, device=DEV, requires_grad=True)
    m3 = base["means3D"] @ _quat_to_R(cr).T + ct
    rq = _quat_mul((cr / cr.norm())[None], base["rotations"])
    ref = GaussianRasterizer(raster_settings=s)(
        means3D=m3, means2D=torch.zeros_like(m3, requires_grad=True), opacities=base["opacities"],
        colors_precomp=None, shs=base["shs"], scales=base["scales"], rotations=rq, cov3D_precomp=None)
    loss_of(ref[0], ref[1], ref[2]).backward()
    # fast path
    cr2 = torch.tensor(q0, device=DEV, requires_grad=True)
    ct2 = torch.tensor(t0, device=DEV, requires_grad=True)
    got = render_relative(s, base["means3D"], base["opacities"], base["shs"], base["scales"], base["rotations"],
                          cr2, ct2)
    loss_of(got[0], got[1], got[2]).backward()
    # images: same math, pose applied to the camera instead of the points -> fp32 rounding only; a
    # Gaussian sitting on a ceil / tile-rectangle boundary may flip (rare), so compare robustly
    assert float((got[3] != ref[3]).float().mean()) < 2e-3
    for a, b, name in ((got[0], ref[0], "color"), (got[1], ref[1], "depth"), (got[2], ref[2], "alpha")):
        err = (a - b).abs()
        assert float(err.median()) < 1e-6 * float(b.abs().max()), name
        assert float((err > 1e-3 * float(b.abs().max())).float().mean()) < 2e-3, name
    assert cr2.grad.shape == (4,) and ct2.grad.shape == (3,)
    gr, gg = cr.grad.cpu().numpy(), cr2.grad.cpu().numpy()
    tr, tg = ct.grad.cpu().numpy(), ct2.grad.cpu().numpy()
    assert np.abs(gg - gr).max() < 1e-3 * np.abs(gr).max() + 1e-9, (gg, gr)
    assert np.abs(tg - tr).max() < 1e-3 * np.abs(tr).max() + 1e-9, (tg, tr)


def test_render_relative_identity_pose_is_the_plain_render():
    from eags_raster.tracking import render_relative
    from gaussian_rasterizer import GaussianRasterizer
    sc = scene.make_scene("ragged", P=2000, seed=8)
    cam = scene.workload_camera("ragged", w2c=scene.rot_xy(1, 2))
    s = _settings_a(cam)
    base = {k: sc[k].to(DEV) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    ref = GaussianRasterizer(raster_settings=s)(
        means3D=base["means3D"], means2D=torch.zeros_like(base["means3D"]), opacities=base["opacities"],
        colors_precomp=None, shs=base["shs"], scales=base["scales"], rotations=base["rotations"], cov3D_precomp=None)
    got = render_relative(s, base["means3D"], base["opacities"], base["shs"], base["scales"], base["rotations"],
                          torch.tensor([2.0, 0, 0, 0], device=DEV), torch.zeros(3, device=DEV))
    assert torch.equal(got[3], ref[3])
    assert float((got[0] - ref[0]).abs().max()) < 1e-5 and float((got[1] - ref[1]).abs().max()) < 1e-5
    with pytest.raises(ValueError, match="sh_degree"):
        render_relative(s._replace(sh_degree=1), base["means3D"], base["opacities"], base["shs"], base["scales"],
                        base["rotations"], torch.tensor([1.0, 0, 0, 0], device=DEV), torch.zeros(3, device=DEV))
