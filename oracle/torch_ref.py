"""Dense float64 PyTorch restatement of the rasterizer forward -- TEST INFRASTRUCTURE ONLY.

Independent of ``eags_oracle.c``: written against SURVEY.md Appendix A.1/A.3 with plain torch ops
so that ``torch.autograd`` yields reference gradients for A.4/A.5 (the "PyTorch-CPU execution of
the same projection/blend math" that BASELINE.json's north_star asks for).  Dense O(pixels x
Gaussians): small scenes only.

Quirk handling: Q1 the min(0.99,.) clamp is straight-through; Q3 p_w has +1e-7;
Q5/Q6 radius, tile rectangles and the skip / termination rules are hard (non-differentiable)
masks; Q4 (frustum clamp) and Q2 are *not* emulated -- use scenes inside 1.3x FoV and a 1e-4
tolerance.  The camera-pose gradient is obtained by differentiating through
``T_CW <- Exp(tau) T_CW`` (reference src/utils/pose_utils.py:61-93) at tau = 0.
"""
from __future__ import annotations

import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792,
      0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
      -0.4570457994644658, 1.445305721320277, -0.5900435899266435]


def _skew(x):
    z = torch.zeros((), dtype=x.dtype)
    return torch.stack([torch.stack([z, -x[2], x[1]]), torch.stack([x[2], z, -x[0]]),
                        torch.stack([-x[1], x[0], z])])


def se3_exp_small(tau):
    """Exp map as the reference evaluates it for |theta| < 1e-5 (pose_utils.py:25-73)."""
    rho, theta = tau[:3], tau[3:]
    Wm = _skew(theta)
    W2 = Wm @ Wm
    eye = torch.eye(3, dtype=tau.dtype)
    R = eye + Wm + 0.5 * W2
    Vm = eye + 0.5 * Wm + (1.0 / 6.0) * W2
    T = torch.eye(4, dtype=tau.dtype)
    T = T.clone()
    T[:3, :3] = R
    T[:3, 3] = Vm @ rho
    return T


def quat_to_R(q):
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.view(-1, 3, 3)


def eval_sh(deg, sh, dirs):
    """sh: (P, M, 3); dirs: (P, 3) unit.  Basis of gaussian_model_utils.py:59-114."""
    res = C0 * sh[:, 0]
    if deg > 0:
        x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
        res = res - C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] +
                   C2[2] * (2 * zz - xx - yy) * sh[:, 6] + C2[3] * xz * sh[:, 7] +
                   C2[4] * (xx - yy) * sh[:, 8])
            if deg > 2:
                res = (res + C3[0] * y * (3 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10] +
                       C3[2] * y * (4 * zz - xx - yy) * sh[:, 11] +
                       C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12] +
                       C3[4] * x * (4 * zz - xx - yy) * sh[:, 13] +
                       C3[5] * z * (xx - yy) * sh[:, 14] + C3[6] * x * (xx - 3 * yy) * sh[:, 15])
    return res


def render(W, H, tanfovx, tanfovy, w2c, proj_raw_T, campos, means3D, scales, rots, opacities,
           shs=None, colors_precomp=None, degree=0, bg=(0.0, 0.0, 0.0), scale_modifier=1.0,
           tau=None):
    """Dense differentiable render.  ``w2c`` (4,4) world->camera, ``proj_raw_T`` = P^T (4,4)
    (the reference's ``projection_matrix``), all float64.  Returns dict(color (3,H,W), depth (H,W),
    alpha_sum (H,W), opacity (H,W) = 1 - T, radii (P,))."""
    dt = torch.float64
    means3D, scales, rots = means3D.to(dt), scales.to(dt), rots.to(dt)
    opacities = opacities.to(dt).reshape(-1)
    w2c = w2c.to(dt)
    if tau is not None:
        w2c = se3_exp_small(tau) @ w2c
    P = means3D.shape[0]
    Rw, tw = w2c[:3, :3], w2c[:3, 3]
    pv = means3D @ Rw.T + tw                                     # (P,3) camera-frame points
    Pm = proj_raw_T.to(dt).T                                     # (4,4) P
    ph = torch.cat([pv, torch.ones(P, 1, dtype=dt)], 1) @ Pm.T   # (P,4)
    pw = 1.0 / (ph[:, 3] + 1e-7)
    ppx, ppy = ph[:, 0] * pw, ph[:, 1] * pw
    fx, fy = W / (2 * tanfovx), H / (2 * tanfovy)
    # Sigma3D
    R = quat_to_R(rots)
    Mx = R * (scale_modifier * scales)[:, None, :]
    Sig = Mx @ Mx.transpose(1, 2)
    # EWA
    tz = pv[:, 2]
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    tx = torch.clamp(pv[:, 0] / tz, -limx, limx) * tz
    ty = torch.clamp(pv[:, 1] / tz, -limy, limy) * tz
    z0 = torch.zeros_like(tz)
    J = torch.stack([fx / tz, z0, -fx * tx / (tz * tz), z0, fy / tz, -fy * ty / (tz * tz)], 1).view(P, 2, 3)
    Mt = J @ Rw
    cov = Mt @ Sig @ Mt.transpose(1, 2)
    a, b, c = cov[:, 0, 0] + 0.3, cov[:, 0, 1], cov[:, 1, 1] + 0.3
    det = a * c - b * b
    conx, cony, conz = c / det, -b / det, a / det
    mid = 0.5 * (a + c)
    lam = mid + torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
    radius = torch.ceil(3.0 * torch.sqrt(lam)).detach()
    pixx = ((ppx + 1.0) * W - 1.0) * 0.5
    pixy = ((ppy + 1.0) * H - 1.0) * 0.5
    gx, gy = (W + 15) // 16, (H + 15) // 16
    with torch.no_grad():
        rminx = torch.clamp(torch.trunc((pixx - radius) / 16), 0, gx)
        rminy = torch.clamp(torch.trunc((pixy - radius) / 16), 0, gy)
        rmaxx = torch.clamp(torch.trunc((pixx + radius + 15) / 16), 0, gx)
        rmaxy = torch.clamp(torch.trunc((pixy + radius + 15) / 16), 0, gy)
        visible = (tz > 0.2) & (det != 0) & ((rmaxx - rminx) * (rmaxy - rminy) > 0)
    if colors_precomp is not None:
        rgb = colors_precomp.to(dt)
    else:
        d = means3D - campos.to(dt)
        d = d / d.norm(dim=1, keepdim=True)
        rgb = torch.clamp_min(eval_sh(degree, shs.to(dt), d) + 0.5, 0.0)
    # depth order (ties by index: stable)
    with torch.no_grad():
        order = torch.argsort(tz.detach(), stable=True)
    ys, xs = torch.meshgrid(torch.arange(H, dtype=dt), torch.arange(W, dtype=dt), indexing="ij")
    pxf, pyf = xs.reshape(-1, 1), ys.reshape(-1, 1)                     # (N,1)
    o = lambda t: t[order]
    dx = o(pixx)[None, :] - pxf
    dy = o(pixy)[None, :] - pyf
    power = -0.5 * (o(conx)[None] * dx * dx + o(conz)[None] * dy * dy) - o(cony)[None] * dx * dy
    G = torch.exp(torch.clamp(power, max=0.0))
    alpha_raw = o(opacities)[None] * G
    alpha = alpha_raw + (torch.clamp(alpha_raw, max=0.99) - alpha_raw).detach()   # Q1
    with torch.no_grad():
        tpx, tpy = torch.floor(pxf / 16), torch.floor(pyf / 16)
        member = (o(visible)[None] & (tpx >= o(rminx)[None]) & (tpx < o(rmaxx)[None]) &
                  (tpy >= o(rminy)[None]) & (tpy < o(rmaxy)[None]))
        keep = member & (power <= 0) & (alpha >= 1.0 / 255.0)
    alpha = torch.where(keep, alpha, torch.zeros_like(alpha))
    one_minus = 1.0 - alpha
    Tincl = torch.cumprod(one_minus, dim=1)                              # T after each entry
    Texcl = torch.cat([torch.ones(Tincl.shape[0], 1, dtype=dt), Tincl[:, :-1]], 1)
    with torch.no_grad():
        stop = (Tincl < 1e-4) & keep
        dead = torch.cumsum(stop.to(torch.int64), dim=1) > 0             # from the first stop on
    w = torch.where(dead, torch.zeros_like(alpha), alpha * Texcl)
    live_T = torch.where(dead, torch.ones_like(one_minus), one_minus)
    Tfinal = torch.prod(torch.where(keep & ~dead, live_T, torch.ones_like(live_T)), dim=1)
    color = w @ o(rgb) + Tfinal[:, None] * torch.tensor(bg, dtype=dt)[None]
    depth = w @ o(tz)
    asum = w.sum(1)
    return dict(color=color.T.reshape(3, H, W), depth=depth.reshape(H, W),
                alpha_sum=asum.reshape(H, W), opacity=(1.0 - Tfinal).reshape(H, W),
                radii=torch.where(visible, radius, torch.zeros_like(radius)).to(torch.int32),
                final_T=Tfinal.reshape(H, W))
