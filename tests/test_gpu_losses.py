"""Fused image-space losses (eags_raster.losses) against a torch restatement of the reference code:
tracker.py:112-128 (tracking loss) and mapper.py:161-165 + losses.py:8-27 (masked L1 means).
Values within 1e-5 relative, gradients within 1e-5 relative (they are signs times a scalar)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _images(H, W, seed, nan_free=True):
    g = torch.Generator(DEV).manual_seed(seed)
    color = torch.rand(3, H, W, device=DEV, generator=g)
    depth = 1 + 3 * torch.rand(1, H, W, device=DEV, generator=g)
    alpha = torch.rand(1, H, W, device=DEV, generator=g)
    gt_c = torch.rand(3, H, W, device=DEV, generator=g)
    gt_d = 1 + 3 * torch.rand(1, H, W, device=DEV, generator=g)
    gt_d[:, : H // 7] = 0                                    # invalid depth rows
    color[:, 3, 5] = gt_c[:, 3, 5]                           # exact zeros: the +1e-8 keeps them "positive"
    depth_mask = (gt_d > 0)
    return color, depth, alpha, gt_c, gt_d, depth_mask


def _ref_tracking(color, depth, alpha, gt_c, gt_d, mask, soft_alpha, mask_color):
    """tracker.py:116-128, line for line in torch."""
    color_loss = torch.abs(color - gt_c) + 1e-8
    depth_loss = (torch.abs(depth - gt_d) + 1e-8) * mask
    if soft_alpha:
        a3 = alpha ** 3
        color_loss = color_loss * a3
        depth_loss = depth_loss * a3
        if mask_color:
            color_loss = color_loss * mask
    else:
        color_loss = color_loss * mask
    return color_loss.sum() / torch.sum(color_loss > 0), depth_loss.sum() / torch.sum(depth_loss > 0)


@pytest.mark.parametrize("soft_alpha,mask_color", [(True, False), (True, True), (False, False)])
@pytest.mark.parametrize("H,W", [(52, 70), (480, 640)])
def test_tracking_losses(soft_alpha, mask_color, H, W):
    from eags_raster.losses import tracking_losses
    color, depth, alpha, gt_c, gt_d, dmask = _images(H, W, 1)
    mask = dmask & (alpha > 0.2)
    leaves = [t.clone().requires_grad_(True) for t in (color, depth, alpha)]
    lc, ld = tracking_losses(leaves[0], leaves[1], leaves[2], gt_c, gt_d, mask, soft_alpha=soft_alpha,
                             mask_invalid_depth_in_color_loss=mask_color)
    (0.7 * lc + 1.3 * ld).backward()
    ref_leaves = [t.clone().double().requires_grad_(True) for t in (color, depth, alpha)]
    rc, rd = _ref_tracking(ref_leaves[0], ref_leaves[1], ref_leaves[2], gt_c.double(), gt_d.double(), mask,
                           soft_alpha, mask_color)
    (0.7 * rc + 1.3 * rd).backward()
    assert abs(float(lc) - float(rc)) <= 1e-5 * abs(float(rc))
    assert abs(float(ld) - float(rd)) <= 1e-5 * abs(float(rd))
    for got, ref, name in zip(leaves, ref_leaves, ("color", "depth", "alpha")):
        if ref.grad is None or (name == "alpha" and not soft_alpha):
            assert got.grad is None or float(got.grad.abs().max()) == 0
            continue
        err = float((got.grad.double() - ref.grad).abs().max())
        assert err <= 1e-5 * float(ref.grad.abs().max()), (name, err)


@pytest.mark.parametrize("use_mask", [False, True])
def test_masked_l1_means(use_mask):
    from eags_raster.losses import masked_l1_means
    H, W = 680, 1200
    color, depth, alpha, gt_c, gt_d, dmask = _images(H, W, 2)
    mask = dmask.squeeze(0) if use_mask else None
    c = color.clone().requires_grad_(True)
    d = depth.clone().requires_grad_(True)
    lc, ld = masked_l1_means(c, d, gt_c, gt_d, mask)
    (lc + ld).backward()
    c2 = color.clone().double().requires_grad_(True)
    d2 = depth.clone().double().requires_grad_(True)
    if use_mask:   # mapper.py:161-165
        rc = torch.abs(c2[:, mask] - gt_c.double()[:, mask]).mean()
        rd = torch.abs(d2[:, mask] - gt_d.double()[0][mask]).mean()
    else:
        rc = torch.abs(c2 - gt_c.double()).mean()
        rd = torch.abs(d2 - gt_d.double()).mean()
    (rc + rd).backward()
    assert abs(float(lc) - float(rc)) <= 1e-5 * abs(float(rc))
    assert abs(float(ld) - float(rd)) <= 1e-5 * abs(float(rd))
    assert float((c.grad.double() - c2.grad).abs().max()) <= 1e-5 * float(c2.grad.abs().max())
    assert float((d.grad.double() - d2.grad).abs().max()) <= 1e-5 * float(d2.grad.abs().max())
    # deterministic (fixed-order reduction): bit-identical reruns
    lc2, ld2 = masked_l1_means(color, depth, gt_c, gt_d, mask)
    assert float(lc2) == float(lc) and float(ld2) == float(ld)


def test_losses_refuse_cpu_tensors():
    from eags_raster.losses import masked_l1_means
    z = torch.zeros(3, 4, 4)
    with pytest.raises(RuntimeError, match="no CPU path"):
        masked_l1_means(z, z[:1], z, z[:1])


# ---- pinned by the reference's own l1_loss, combined as its call sites combine it -------------------
# tests/golden/losses_reference.npz (make_golden_losses.py: tracker.py:112-126, mapper.py:160-165, fp32 CPU
# torch + autograd).  Values 1e-5 relative; gradients 1e-5 relative to the largest reference gradient.
def _golden():
    import os
    import numpy as np
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "losses_reference.npz"))


@pytest.mark.parametrize("tag,soft_alpha,mask_color", [("soft", True, False), ("soft_masked", True, True),
                                                       ("hard", False, False)])
def test_tracking_losses_against_reference_vectors(tag, soft_alpha, mask_color):
    from eags_raster.losses import tracking_losses
    g = _golden()
    t = {k: torch.from_numpy(g[k]).to(DEV) for k in ("color", "depth", "alpha", "gt_color", "gt_depth", "tracking_mask")}
    c, d, a = (t[k].clone().requires_grad_(True) for k in ("color", "depth", "alpha"))
    lc, ld = tracking_losses(c, d, a, t["gt_color"], t["gt_depth"], t["tracking_mask"], soft_alpha=soft_alpha,
                             mask_invalid_depth_in_color_loss=mask_color)
    (0.7 * lc + 1.3 * ld).backward()
    rc, rd = float(g[f"track_{tag}_color_loss"]), float(g[f"track_{tag}_depth_loss"])
    assert abs(float(lc) - rc) <= 1e-5 * abs(rc) and abs(float(ld) - rd) <= 1e-5 * abs(rd)
    for leaf, key in ((c, "dcolor"), (d, "ddepth"), (a, "dalpha")):
        ref = torch.from_numpy(g[f"track_{tag}_{key}"]).to(DEV)
        got = leaf.grad if leaf.grad is not None else torch.zeros_like(ref)
        assert float((got - ref).abs().max()) <= 1e-5 * float(ref.abs().max()) + 1e-12, key


def test_masked_l1_means_against_reference_vectors():
    from eags_raster.losses import masked_l1_means
    g = _golden()
    t = {k: torch.from_numpy(g[k]).to(DEV) for k in ("color", "depth", "gt_color", "gt_depth", "map_mask")}
    c, d = t["color"].clone().requires_grad_(True), t["depth"].clone().requires_grad_(True)
    lc, ld = masked_l1_means(c, d, t["gt_color"], t["gt_depth"], t["map_mask"])
    (lc + ld).backward()
    assert abs(float(lc) - float(g["map_color_l1"])) <= 1e-5 * float(g["map_color_l1"])
    assert abs(float(ld) - float(g["map_depth_l1"])) <= 1e-5 * float(g["map_depth_l1"])
    for leaf, key in ((c, "map_dcolor"), (d, "map_ddepth")):
        ref = torch.from_numpy(g[key]).to(DEV)
        assert float((leaf.grad - ref).abs().max()) <= 1e-5 * float(ref.abs().max()), key


def test_registration_loss_matches_the_references_get_loss_tracking_rgbd():
    """eags_raster.losses.registration_loss against golden vectors produced by the reference's own
    get_loss_tracking_rgbd (src/gsr/loss.py:66-95) and its autograd gradients
    (tests/golden/make_golden_registration.py)."""
    import os
    import numpy as np
    from eags_raster.losses import registration_loss
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "registration_reference.npz"))
    dev = torch.device("cuda")
    for case in range(3):
        t = {k: torch.from_numpy(g[f"loss{case}_{k}"]).to(dev) for k in
             ("image", "depth", "opacity", "gt", "gt_depth", "grad_mask", "loss", "g_image", "g_depth", "g_opacity")}
        image, depth, opacity = (t[k].clone().requires_grad_(True) for k in ("image", "depth", "opacity"))
        rgb_mask = (t["gt"].sum(dim=0) > 0.01) & (t["grad_mask"][0] > 0)
        depth_mask = t["gt_depth"] > 0.01
        loss = registration_loss(image, depth, opacity, t["gt"], t["gt_depth"][None], rgb_mask, depth_mask)
        loss.backward()
        assert abs(float(loss) - float(t["loss"])) <= 1e-6 * abs(float(t["loss"]))
        for got, want in ((image.grad, t["g_image"]), (depth.grad, t["g_depth"]), (opacity.grad, t["g_opacity"])):
            assert float((got - want).abs().max()) <= 1e-6 * float(want.abs().max()) + 1e-12
