"""Golden vectors for the fused SSIM, produced by the REFERENCE's own implementation
(/root/reference/src/entities/losses.py::ssim, CPU) and its autograd gradient.  Run in the build
container (the reference is not present on the GPU box):  python tests/golden/make_golden_ssim.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from src.entities.losses import l1_loss, ssim  # noqa: E402

out = {}
for name, (C, H, W), seed in [("ragged", (3, 37, 53), 0), ("square", (3, 64, 64), 1), ("gray", (1, 20, 45), 2)]:
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(C, H, W, generator=g)
    img = (gt + 0.15 * torch.randn(C, H, W, generator=g)).clamp(0, 1).requires_grad_(True)
    val = ssim(img, gt)
    val.backward()
    out[f"{name}_img"] = img.detach().numpy()
    out[f"{name}_gt"] = gt.numpy()
    out[f"{name}_ssim"] = np.float32(val.item())
    out[f"{name}_grad"] = img.grad.numpy().copy()
    # the mapping colour loss of mapper.py:162-163 with lambda_dssim = 0.2 and an all-true mask
    img2 = img.detach().clone().requires_grad_(True)
    mask = torch.ones(H, W, dtype=torch.bool)
    loss = 0.8 * l1_loss(img2[:, mask], gt[:, mask]) + 0.2 * (1.0 - ssim(img2, gt))
    loss.backward()
    out[f"{name}_maploss"] = np.float32(loss.item())
    out[f"{name}_maploss_grad"] = img2.grad.numpy().copy()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ssim_reference.npz"), **out)
print("wrote", len(out), "arrays")

# isotropic regulariser (src/entities/losses.py:115-126), value and gradient for a 2x upstream gradient
from src.entities.losses import isotropic_loss  # noqa: E402

g = torch.Generator().manual_seed(9)
s = torch.exp(torch.randn(5001, 3, generator=g) * 0.7 - 3).requires_grad_(True)
with torch.no_grad():
    s[:50] = s[:50, :1]          # exactly isotropic rows: |0| has zero gradient
v = isotropic_loss(s)
(2.0 * v).backward()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "isotropic_reference.npz"),
                    scaling=s.detach().numpy(), loss=np.float32(v.item()), grad=s.grad.numpy())
print("wrote isotropic_reference.npz")
