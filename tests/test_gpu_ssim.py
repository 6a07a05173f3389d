"""GPU parity of the fused SSIM kernels (csrc/ssim.cu through the C ABI, eags_raster.losses.ssim /
mapping_color_loss) against the REFERENCE's own outputs (golden vectors of src/entities/losses.py::ssim and of
the mapping colour loss of src/entities/mapper.py:162-163) and, at the benchmark image size, against the
float64 oracle.  Tolerances: value 2e-6 absolute (fp32 vs the reference's fp32 conv2d), gradient 1e-4 of its
maximum."""
import os

import numpy as np
import pytest
import torch

from oracle import ssim as oss

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ssim_reference.npz"))


@pytest.mark.parametrize("name", ["ragged", "square", "gray"])
def test_fused_ssim_matches_reference_golden(name):
    from eags_raster.losses import mapping_color_loss, ssim
    img = torch.from_numpy(G[f"{name}_img"]).cuda().requires_grad_(True)
    gt = torch.from_numpy(G[f"{name}_gt"]).cuda()
    val = ssim(img, gt)
    assert val.shape == () and abs(float(val) - float(G[f"{name}_ssim"])) < 2e-6
    (3.0 * val).backward()                      # a non-unit upstream gradient, read on the device
    ref = 3.0 * G[f"{name}_grad"]
    assert np.abs(img.grad.cpu().numpy() - ref).max() < 1e-4 * np.abs(ref).max()
    if img.shape[0] != 3:
        return                                  # the mapping colour loss is RGB (mapper.py:162)
    img2 = torch.from_numpy(G[f"{name}_img"]).cuda().requires_grad_(True)
    mask = torch.ones(img2.shape[-2:], dtype=torch.bool, device="cuda")
    loss = mapping_color_loss(img2, gt, mask, 0.2)
    assert abs(float(loss) - float(G[f"{name}_maploss"])) < 2e-6
    loss.backward()
    ref = G[f"{name}_maploss_grad"]
    assert np.abs(img2.grad.cpu().numpy() - ref).max() < 1e-4 * np.abs(ref).max()


def test_fused_ssim_full_size_against_oracle():
    from eags_raster.losses import ssim
    g = torch.Generator().manual_seed(5)
    gt = torch.rand(3, 480, 640, generator=g)
    img = (gt + 0.1 * torch.randn(3, 480, 640, generator=g)).clamp(0, 1)
    val_o, grad_o = oss.ssim_and_grad(img.numpy(), gt.numpy())
    x = img.cuda().requires_grad_(True)
    val = ssim(x, gt.cuda())
    val.backward()
    assert abs(float(val) - val_o) < 2e-6
    assert np.abs(x.grad.cpu().numpy() - grad_o).max() < 1e-4 * np.abs(grad_o).max()
    assert float(ssim(gt.cuda(), gt.cuda())) == pytest.approx(1.0, abs=1e-6)


def test_fused_ssim_argument_errors():
    from eags_raster.losses import ssim
    a = torch.rand(3, 16, 16, device="cuda")
    with pytest.raises(NotImplementedError):
        ssim(a, a, window_size=7)
    with pytest.raises(ValueError):
        ssim(a, a[:, :8])
    with pytest.raises(RuntimeError, match="CUDA"):
        ssim(a.cpu(), a.cpu())


def test_fused_isotropic_loss_matches_reference_golden():
    """src/entities/losses.py:115-126 through the reference's own code (tests/golden/isotropic_reference.npz)."""
    from eags_raster.losses import isotropic_loss
    R = np.load(os.path.join(os.path.dirname(__file__), "golden", "isotropic_reference.npz"))
    s = torch.from_numpy(R["scaling"]).cuda().requires_grad_(True)
    v = isotropic_loss(s)
    assert abs(float(v) - float(R["loss"])) <= 2e-6 * float(R["loss"])
    (2.0 * v).backward()
    assert np.abs(s.grad.cpu().numpy() - R["grad"]).max() <= 1e-6 * np.abs(R["grad"]).max()
    with pytest.raises(ValueError):
        isotropic_loss(torch.zeros(0, 3, device="cuda"))


ML = np.load(os.path.join(os.path.dirname(__file__), "golden", "mapping_loss_reference.npz"))


@pytest.mark.parametrize("name", ["ragged", "tile", "nomask", "empty"])
def test_fused_mapping_losses_match_the_references_mapper_composition(name):
    """eags_raster.losses.mapping_losses (one forward + one backward kernel) against the reference's own l1_loss
    and ssim composed as src/entities/mapper.py:161-165 (tests/golden/make_golden_mapping_loss.py): partial
    mask, exact ties, no mask, an empty mask (nan losses, SSIM-only gradient), upstream weights 2 and 3."""
    from eags_raster.losses import mapping_losses
    img = torch.from_numpy(ML[f"{name}_img"]).cuda().requires_grad_(True)
    dep = torch.from_numpy(ML[f"{name}_depth"]).cuda().requires_grad_(True)
    gt, gt_d = torch.from_numpy(ML[f"{name}_gt"]).cuda(), torch.from_numpy(ML[f"{name}_gt_depth"]).cuda()
    mask = None if name == "nomask" else torch.from_numpy(ML[f"{name}_mask"]).cuda()
    c, d = mapping_losses(img, dep, gt, gt_d, mask, 0.2)
    if name == "empty":
        assert np.isnan(float(c)) and np.isnan(float(d)) and np.isnan(float(ML["empty_color_loss"]))
        # the reference's empty means carry no gradient: what is left is the SSIM term
        (2.0 * c + 3.0 * d).backward()
    else:
        assert abs(float(c) - float(ML[f"{name}_color_loss"])) < 2e-6
        assert abs(float(d) - float(ML[f"{name}_depth_loss"])) <= 2e-6 * float(ML[f"{name}_depth_loss"])
        (2.0 * c + 3.0 * d).backward()
    ref_i, ref_d = ML[f"{name}_grad_img"], ML[f"{name}_grad_depth"]
    assert np.abs(img.grad.cpu().numpy() - ref_i).max() < 1e-4 * np.abs(ref_i).max()
    assert np.abs(dep.grad.cpu().numpy() - ref_d).max() <= 1e-6 * max(np.abs(ref_d).max(), 1e-30)


def test_fused_mapping_losses_equal_the_separate_kernels_and_are_reproducible():
    """At the benchmark's image size (1200x680, a ragged tile edge): the fused pair against
    mapping_color_loss + masked_l1_means (each pinned above / in test_gpu_losses.py); two runs give the same bits."""
    from eags_raster.losses import mapping_color_loss, mapping_losses, masked_l1_means
    g = torch.Generator().manual_seed(11)
    H, W = 680, 1200
    gt = torch.rand(3, H, W, generator=g).cuda()
    img0 = (gt + 0.1 * torch.randn(3, H, W, generator=g).cuda()).clamp(0, 1)
    gt_d = (1.0 + 3.0 * torch.rand(1, H, W, generator=g)).cuda()
    dep0 = gt_d + 0.2 * torch.randn(1, H, W, generator=g).cuda()
    mask = (torch.rand(H, W, generator=g) > 0.3).cuda()
    runs = []
    for _ in range(2):
        img, dep = img0.clone().requires_grad_(True), dep0.clone().requires_grad_(True)
        c, d = mapping_losses(img, dep, gt, gt_d, mask, 0.2)
        (c + d).backward()
        runs.append((c.item(), d.item(), img.grad.clone(), dep.grad.clone()))
    assert runs[0][0] == runs[1][0] and runs[0][1] == runs[1][1]
    assert torch.equal(runs[0][2], runs[1][2]) and torch.equal(runs[0][3], runs[1][3])
    img, dep = img0.clone().requires_grad_(True), dep0.clone().requires_grad_(True)
    _, d2 = masked_l1_means(img, dep, gt, gt_d, mask)
    c2 = mapping_color_loss(img, gt, mask, 0.2)
    (c2 + d2).backward()
    assert abs(runs[0][0] - c2.item()) < 1e-6 and abs(runs[0][1] - d2.item()) <= 1e-6 * d2.item()
    gi, gd = runs[0][2], runs[0][3]
    assert float((gi - img.grad).abs().max()) <= 1e-5 * float(img.grad.abs().max())
    assert float((gd - dep.grad).abs().max()) <= 1e-6 * float(dep.grad.abs().max())
    with pytest.raises(ValueError):
        mapping_losses(img0, dep0[:, :8], gt, gt_d, None, 0.2)
    with pytest.raises(RuntimeError, match="CUDA"):
        mapping_losses(img0.cpu(), dep0.cpu(), gt.cpu(), gt_d.cpu(), None, 0.2)


def test_fused_mapping_losses_with_the_mappers_own_mask():
    """mask="valid_depth": (gt_depth > 0) & ~isnan(depth) (mapper.py:159) formed inside the kernels; holes in
    gt_depth and nan pixels in the rendered depth, against the reference's expression (golden `vd_*`), and equal
    to passing that mask as a tensor."""
    from eags_raster.losses import mapping_losses
    gt, gt_d = torch.from_numpy(ML["vd_gt"]).cuda(), torch.from_numpy(ML["vd_gt_depth"]).cuda()
    res = []
    for mask in ("valid_depth", torch.from_numpy(ML["vd_mask"]).cuda()):
        img = torch.from_numpy(ML["vd_img"]).cuda().requires_grad_(True)
        dep = torch.from_numpy(ML["vd_depth"]).cuda().requires_grad_(True)
        c, d = mapping_losses(img, dep, gt, gt_d, mask, 0.2)
        assert abs(float(c) - float(ML["vd_color_loss"])) < 2e-6
        assert abs(float(d) - float(ML["vd_depth_loss"])) <= 2e-6 * float(ML["vd_depth_loss"])
        (2.0 * c + 3.0 * d).backward()
        ref_i, ref_d = ML["vd_grad_img"], ML["vd_grad_depth"]
        assert not torch.isnan(img.grad).any() and not torch.isnan(dep.grad).any()
        assert np.abs(img.grad.cpu().numpy() - ref_i).max() < 1e-4 * np.abs(ref_i).max()
        assert np.abs(dep.grad.cpu().numpy() - ref_d).max() <= 1e-6 * np.abs(ref_d).max()
        res.append((c.item(), d.item(), img.grad.clone(), dep.grad.clone()))
    assert res[0][0] == res[1][0] and res[0][1] == res[1][1]
    assert torch.equal(res[0][2], res[1][2]) and torch.equal(res[0][3], res[1][3])
    with pytest.raises(ValueError):
        mapping_losses(res[0][2], res[0][3], gt, gt_d, "depth", 0.2)
