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
