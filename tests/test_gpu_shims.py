"""The two operator shims driven the way the reference's call sites drive them, with autograd:

  * variant A as src/utils/utils.py:105-117,149-184 (settings by keyword, viewmatrix passed as a
    transposed *view*, means2D zeros with retain_grad, 4-tuple) inside a tracking-style step
    (src/entities/tracker.py:85-97: all means transformed by an optimisable relative pose, rotations
    left-multiplied by the camera quaternion, pose gradient through dL/dmeans3D and dL/drotations) and
    a mapping-style step (src/entities/mapper.py:152-170: activations, SH frozen);
  * variant B as src/gsr/renderer.py:59-73,131-142 (13-field settings, theta / rho parameters,
    5-tuple) with the pose gradient checked against the oracle's closed form.

Gradients are compared with the oracle's analytic backward chained through float64 torch autograd
of the surrounding PyTorch ops.
"""
import numpy as np
import pytest
import torch

import oracle
from eags_raster import scene
from helpers import nonfragile, oracle_forward, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _settings_a(cam):
    from gaussian_rasterizer import GaussianRasterizationSettings
    w2c = cam.w2c.to(DEV)
    viewmatrix = w2c.transpose(0, 1)                      # a non-contiguous view, as utils.py:97
    return GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=torch.tensor([0, 0, 0], dtype=torch.float, device=DEV), scale_modifier=1.0,
        viewmatrix=viewmatrix, projmatrix=cam.projmatrix.to(DEV), sh_degree=0,
        campos=torch.inverse(w2c)[:3, 3], prefiltered=False, debug=False)


def _quat_mul(q1, q2):
    w1, x1, y1, z1 = q1.unbind(-1)
    w2, x2, y2, z2 = q2.unbind(-1)
    return torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                        w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], -1)


def _quat_to_R(q):
    q = q / q.norm()
    r, x, y, z = q
    return torch.stack([
        torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)]),
        torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)]),
        torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)])])


def _oracle_grads(cam, tensors, dC_fn):
    """Oracle forward + backward on numpy copies; dC_fn(o) -> (dC, dD, dA) upstream gradients."""
    sc = {k: v.detach().cpu().float() for k, v in tensors.items()}
    o = oracle_forward(cam, sc)
    dC, dD, dA = dC_fn(o)
    ok = nonfragile(o)
    for g in (dC, dD, dA):
        if g is not None:
            g[:, ~ok] = 0
    return o, ok, oracle.backward(o, dC, dD, dA, want_tau=True)


def test_variant_a_tracking_style_pose_gradient():
    from gaussian_rasterizer import GaussianRasterizer
    sc = scene.make_scene("ragged", P=2500, seed=7)
    cam = scene.workload_camera("ragged")
    s = _settings_a(cam)
    base = {k: sc[k].to(DEV) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    # optimisable relative pose (tracker.py:85-95)
    cam_rot = torch.tensor([1.0, 0.004, -0.003, 0.002], device=DEV, requires_grad=True)
    cam_trans = torch.tensor([0.01, -0.005, 0.008], device=DEV, requires_grad=True)

    def transformed(cr, ct, means, rots):
        R = _quat_to_R(cr)
        m = means @ R.T + ct
        q = _quat_mul((cr / cr.norm())[None], rots)
        return m, q

    m3, rq = transformed(cam_rot, cam_trans, base["means3D"], base["rotations"])
    means2D = torch.zeros_like(m3, dtype=m3.dtype, requires_grad=True, device=DEV)
    means2D.retain_grad()
    color, depth, alpha, radii = GaussianRasterizer(raster_settings=s)(
        means3D=m3, means2D=means2D, opacities=base["opacities"], colors_precomp=None, shs=base["shs"],
        scales=base["scales"], rotations=rq, cov3D_precomp=None)
    assert color.shape == (3, cam.H, cam.W) and depth.shape == (1, cam.H, cam.W) and alpha.shape == depth.shape
    assert radii.dtype == torch.int32 and radii.shape == (2500,)
    gt_c, gt_d = sc["gt_color"].to(DEV), sc["gt_depth"].to(DEV)
    wc = torch.randn(3, cam.H, cam.W, device=DEV, generator=torch.Generator(DEV).manual_seed(1))
    wd = torch.randn(1, cam.H, cam.W, device=DEV, generator=torch.Generator(DEV).manual_seed(2))
    # fragile pixels are masked out of the loss on both sides
    o, ok, ref = _oracle_grads(cam, dict(base, means3D=m3, rotations=rq),
                               lambda o: (wc.cpu().numpy().copy(), wd.cpu().numpy().copy(),
                                          (0.5 * wd).cpu().numpy().copy()))
    okt = torch.from_numpy(ok).to(DEV)
    loss = ((color * wc)[:, okt]).sum() + ((depth * wd)[:, okt]).sum() + ((alpha * 0.5 * wd)[:, okt]).sum()
    loss.backward()
    # chain the oracle's dL/dmeans3D, dL/drotations through the same transform in float64
    cr = cam_rot.detach().double().cpu().requires_grad_(True)
    ct = cam_trans.detach().double().cpu().requires_grad_(True)
    m64, q64 = transformed(cr, ct, base["means3D"].double().cpu(), base["rotations"].double().cpu())
    (m64 * torch.from_numpy(ref["dL_dmeans3D"]).double()).sum().add(
        (q64 * torch.from_numpy(ref["dL_drots"]).double()).sum()).backward()
    assert rel_err(cam_rot.grad.cpu().numpy(), cr.grad.numpy()) < 1e-3
    assert rel_err(cam_trans.grad.cpu().numpy(), ct.grad.numpy()) < 1e-3
    # the screen-space gradient lands in means2D.grad (utils.py:157-159)
    assert means2D.grad is not None and rel_err(means2D.grad[:, :2].cpu().numpy(), ref["dL_dmean2D"]) < 1e-3


def test_variant_a_mapping_style_step():
    from gaussian_rasterizer import GaussianRasterizer
    sc = scene.make_scene("ragged", P=2000, seed=9, M=16)
    cam = scene.workload_camera("ragged", w2c=scene.rot_xy(2, 1))
    s = _settings_a(cam)
    # raw parameters + activations (gaussian_model.py:103-122); SH features frozen (mapper.py:358-359)
    xyz = sc["means3D"].to(DEV).requires_grad_(True)
    scaling = torch.log(sc["scales"]).to(DEV).requires_grad_(True)
    rotation = (sc["rotations"] * 1.7).to(DEV).requires_grad_(True)
    opacity = torch.logit(sc["opacities"].clamp(1e-4, 1 - 1e-4)).to(DEV).requires_grad_(True)
    shs = sc["shs"].to(DEV)
    act = lambda sc_, ro_, op_: (torch.exp(sc_), torch.nn.functional.normalize(ro_), torch.sigmoid(op_))
    S, Rn, O = act(scaling, rotation, opacity)
    means2D = torch.zeros_like(xyz, requires_grad=True)
    color, depth, alpha, radii = GaussianRasterizer(raster_settings=s)(
        means3D=xyz, means2D=means2D, opacities=O, colors_precomp=None, shs=shs, scales=S, rotations=Rn,
        cov3D_precomp=None)
    wc = torch.randn(3, cam.H, cam.W, device=DEV, generator=torch.Generator(DEV).manual_seed(4))
    wd = torch.randn(1, cam.H, cam.W, device=DEV, generator=torch.Generator(DEV).manual_seed(5))
    o, ok, ref = _oracle_grads(cam, dict(means3D=xyz, scales=S, rotations=Rn, opacities=O, shs=shs),
                               lambda o: (wc.cpu().numpy().copy(), wd.cpu().numpy().copy(), None))
    okt = torch.from_numpy(ok).to(DEV)
    (((color * wc)[:, okt]).sum() + ((depth * wd)[:, okt]).sum()).backward()
    assert shs.grad is None
    sc64 = scaling.detach().double().cpu().requires_grad_(True)
    ro64 = rotation.detach().double().cpu().requires_grad_(True)
    op64 = opacity.detach().double().cpu().requires_grad_(True)
    S64, R64, O64 = act(sc64, ro64, op64)
    ((S64 * torch.from_numpy(ref["dL_dscales"]).double()).sum() + (R64 * torch.from_numpy(ref["dL_drots"]).double()).sum()
     + (O64[:, 0] * torch.from_numpy(ref["dL_dopacity"]).double()).sum()).backward()
    assert rel_err(xyz.grad.cpu().numpy(), ref["dL_dmeans3D"]) < 1e-3
    assert rel_err(scaling.grad.cpu().numpy(), sc64.grad.numpy()) < 1e-3
    assert rel_err(rotation.grad.cpu().numpy(), ro64.grad.numpy()) < 1e-3
    assert rel_err(opacity.grad.cpu().numpy(), op64.grad.numpy()) < 1e-3
    # shapes may change between calls (prune: mapper.py:201-223)
    keep = radii > 0
    out2 = GaussianRasterizer(raster_settings=s)(
        means3D=xyz[keep], means2D=torch.zeros_like(xyz[keep]), opacities=O.detach()[keep], colors_precomp=None,
        shs=shs[keep], scales=S.detach()[keep], rotations=Rn.detach()[keep], cov3D_precomp=None)
    assert torch.equal(out2[0], color.detach()) and out2[3].shape[0] == int(keep.sum())


def test_variant_b_pose_gradient_through_theta_rho():
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    sc = scene.make_scene("ragged", P=2500, seed=12)
    cam = scene.workload_camera("ragged", w2c=scene.rot_xy(-3, 2))
    s = GaussianRasterizationSettings(
        image_height=int(cam.H), image_width=int(cam.W), tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=torch.zeros(3, device=DEV), scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(DEV),
        projmatrix=cam.projmatrix.to(DEV), projmatrix_raw=cam.projmatrix_raw.to(DEV), sh_degree=0,
        campos=cam.campos.to(DEV), prefiltered=False, debug=False)
    t = {k: sc[k].to(DEV).requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    theta = torch.nn.Parameter(torch.zeros(3, device=DEV))      # camera.py:52-57
    rho = torch.nn.Parameter(torch.zeros(3, device=DEV))
    screenspace = torch.zeros_like(t["means3D"], requires_grad=True)
    screenspace.retain_grad()
    color, radii, depth, opacity, n_touched = GaussianRasterizer(raster_settings=s)(
        means3D=t["means3D"], means2D=screenspace, shs=t["shs"], colors_precomp=None, opacities=t["opacities"],
        scales=t["scales"], rotations=t["rotations"], cov3D_precomp=None, theta=theta, rho=rho)
    assert opacity.shape == (1, cam.H, cam.W) and n_touched.shape == (2500,) and n_touched.dtype == torch.int32
    wc = torch.randn(3, cam.H, cam.W, device=DEV, generator=torch.Generator(DEV).manual_seed(6))
    wd = torch.randn(1, cam.H, cam.W, device=DEV, generator=torch.Generator(DEV).manual_seed(7))
    o, ok, ref = _oracle_grads(cam, {k: v for k, v in t.items()},
                               lambda o: (wc.cpu().numpy().copy(), wd.cpu().numpy().copy(), None))
    okt = torch.from_numpy(ok).to(DEV)
    (((color * wc)[:, okt]).sum() + ((depth * wd)[:, okt]).sum()).backward()
    assert theta.grad.shape == (3,) and rho.grad.shape == (3,)
    assert rel_err(rho.grad.cpu().numpy(), ref["dL_dtau"][:3]) < 1e-3
    assert rel_err(theta.grad.cpu().numpy(), ref["dL_dtau"][3:]) < 1e-3
    assert rel_err(t["means3D"].grad.cpu().numpy(), ref["dL_dmeans3D"]) < 1e-3
    assert rel_err(t["shs"].grad.cpu().numpy(), ref["dL_dshs"]) < 1e-3
    assert screenspace.grad is not None
    vis = radii > 0
    assert int(n_touched[~vis].sum()) == 0
    # registration style (src/gsr/solver.py): only the pose is optimised -> pose-only backward, no
    # per-Gaussian gradient is produced; theta / rho gradients are the same
    th2 = torch.nn.Parameter(torch.zeros(3, device=DEV))
    rh2 = torch.nn.Parameter(torch.zeros(3, device=DEV))
    frozen = {k: v.detach() for k, v in t.items()}
    c2, _, d2, _, _ = GaussianRasterizer(raster_settings=s)(
        means3D=frozen["means3D"], means2D=torch.zeros_like(frozen["means3D"]), shs=frozen["shs"], colors_precomp=None,
        opacities=frozen["opacities"], scales=frozen["scales"], rotations=frozen["rotations"], cov3D_precomp=None,
        theta=th2, rho=rh2)
    (((c2 * wc)[:, okt]).sum() + ((d2 * wd)[:, okt]).sum()).backward()
    assert rel_err(th2.grad.cpu().numpy(), theta.grad.cpu().numpy()) < 1e-5
    assert rel_err(rh2.grad.cpu().numpy(), rho.grad.cpu().numpy()) < 1e-5
    # only ONE of the two pose slots given (backward must return None for the other, not a tensor)
    th3 = torch.nn.Parameter(torch.zeros(3, device=DEV))
    c3, _, d3, _, _ = GaussianRasterizer(raster_settings=s)(
        means3D=frozen["means3D"], means2D=torch.zeros_like(frozen["means3D"]), shs=frozen["shs"], colors_precomp=None,
        opacities=frozen["opacities"], scales=frozen["scales"], rotations=frozen["rotations"], cov3D_precomp=None,
        theta=th3, rho=None)
    (((c3 * wc)[:, okt]).sum() + ((d3 * wd)[:, okt]).sum()).backward()
    assert rel_err(th3.grad.cpu().numpy(), theta.grad.cpu().numpy()) < 1e-5



def test_grad_arena_writes_into_fused_bucket():
    """eags_raster.grad_arena: the backward kernel writes the per-Gaussian gradients straight into
    slices of one flat buffer (the all-reduce bucket of sharded mapping) -- same values, no copies."""
    import eags_raster
    from gaussian_rasterizer import GaussianRasterizer
    sc = scene.make_scene("ragged", P=1800, seed=21)
    cam = scene.workload_camera("ragged")
    s = _settings_a(cam)
    keys = ("means3D", "scales", "rotations", "opacities")
    sizes = {"means3D": 3, "scales": 3, "rotations": 4, "opacities": 1}
    P = 1800

    def run(arena):
        leaf = {k: sc[k].to(DEV).requires_grad_(True) for k in keys}
        out = GaussianRasterizer(raster_settings=s)(
            means3D=leaf["means3D"], means2D=torch.zeros_like(leaf["means3D"], requires_grad=True),
            opacities=leaf["opacities"], colors_precomp=None, shs=sc["shs"].to(DEV), scales=leaf["scales"],
            rotations=leaf["rotations"], cov3D_precomp=None)
        loss = out[0].sum() + 0.3 * out[1].sum()
        if arena is None:
            loss.backward()
        else:
            with eags_raster.grad_arena(**arena):
                loss.backward()
        return leaf

    ref = run(None)
    flat = torch.full((P * 11,), float("nan"), device=DEV)
    views, off = {}, 0
    for k in keys:
        views[k] = flat[off:off + P * sizes[k]].view(P, sizes[k])
        off += P * sizes[k]
    got = run(views)
    assert torch.isfinite(flat).all()
    off = 0
    for k in keys:
        assert rel_err(got[k].grad.cpu().numpy(), ref[k].grad.cpu().numpy()) < 1e-5
        assert torch.equal(flat[off:off + P * sizes[k]].view(P, sizes[k]), got[k].grad)
        assert got[k].grad.untyped_storage().data_ptr() == flat.untyped_storage().data_ptr()   # stolen view, no copy
        off += P * sizes[k]
    with pytest.raises(ValueError, match="grad_arena"):
        run(dict(means3D=torch.zeros(5, 3, device=DEV)))
