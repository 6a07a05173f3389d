"""Golden vectors for the fused per-view mapping loss (eags_raster.losses.mapping_losses), produced by the
REFERENCE's own l1_loss and ssim (/root/reference/src/entities/losses.py) composed exactly as
src/entities/mapper.py:161-165 does, with a partial mask, and autograd for the gradients of
``2 * color_loss + 3 * depth_loss``.  Run in the build container (the reference is not present on the GPU box):
    python tests/golden/make_golden_mapping_loss.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from src.entities.losses import l1_loss, ssim  # noqa: E402

LAMBDA = 0.2
out = {}
for name, (H, W), seed, masked in [("ragged", (37, 53), 0, True), ("tile", (64, 96), 1, True),
                                   ("nomask", (45, 70), 2, False), ("empty", (33, 34), 3, True)]:
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(3, H, W, generator=g)
    img = (gt + 0.15 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
    gt_d = 1.0 + 3.0 * torch.rand(1, H, W, generator=g)
    dep = gt_d + 0.2 * torch.randn(1, H, W, generator=g)
    with torch.no_grad():                      # exact ties: |0| has zero gradient
        img[:, :3, :5] = gt[:, :3, :5]
        dep[:, 4:6, :7] = gt_d[:, 4:6, :7]
    mask = (gt_d[0] > 1.6) if masked else torch.ones(H, W, dtype=torch.bool)
    if name == "empty":
        mask = torch.zeros(H, W, dtype=torch.bool)
    img.requires_grad_(True)
    dep.requires_grad_(True)
    # mapper.py:161-165
    color_loss = (1.0 - LAMBDA) * l1_loss(img[:, mask], gt[:, mask]) + LAMBDA * (1.0 - ssim(img, gt))
    depth_loss = l1_loss(dep[:, mask], gt_d[0][mask])      # gt_depth is (H,W) in the mapper
    if name == "empty":
        (2.0 * LAMBDA * (1.0 - ssim(img, gt))).backward()      # the empty means are nan and carry no gradient
    else:
        (2.0 * color_loss + 3.0 * depth_loss).backward()
    out[f"{name}_img"] = img.detach().numpy()
    out[f"{name}_gt"] = gt.numpy()
    out[f"{name}_depth"] = dep.detach().numpy()
    out[f"{name}_gt_depth"] = gt_d.numpy()
    out[f"{name}_mask"] = mask.numpy()
    out[f"{name}_color_loss"] = np.float32(color_loss.item())
    out[f"{name}_depth_loss"] = np.float32(depth_loss.item())
    out[f"{name}_grad_img"] = img.grad.numpy().copy()
    out[f"{name}_grad_depth"] = dep.grad.numpy().copy() if dep.grad is not None else np.zeros((1, H, W), np.float32)
# the mapper's own mask (mapper.py:159): gt_depth has holes (0), the rendered depth has nan pixels
H, W = 41, 59
g = torch.Generator().manual_seed(7)
gt = torch.rand(3, H, W, generator=g)
img = (gt + 0.15 * torch.randn(3, H, W, generator=g)).clamp(0, 1).requires_grad_(True)
gt_d = 1.0 + 3.0 * torch.rand(H, W, generator=g)
gt_d[torch.rand(H, W, generator=g) < 0.2] = 0.0
dep0 = (gt_d + 0.2 * torch.randn(H, W, generator=g)).unsqueeze(0)
dep0[0][torch.rand(H, W, generator=g) < 0.1] = float("nan")
dep = dep0.clone().requires_grad_(True)
mask = (gt_d > 0) & (~torch.isnan(dep)).squeeze(0)
color_loss = (1.0 - LAMBDA) * l1_loss(img[:, mask], gt[:, mask]) + LAMBDA * (1.0 - ssim(img, gt))
depth_loss = l1_loss(dep[:, mask], gt_d[mask])
(2.0 * color_loss + 3.0 * depth_loss).backward()
out.update(vd_img=img.detach().numpy(), vd_gt=gt.numpy(), vd_depth=dep.detach().numpy(), vd_gt_depth=gt_d.unsqueeze(0).numpy(),
           vd_mask=mask.numpy(), vd_color_loss=np.float32(color_loss.item()), vd_depth_loss=np.float32(depth_loss.item()),
           vd_grad_img=img.grad.numpy().copy(), vd_grad_depth=dep.grad.numpy().copy())
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "mapping_loss_reference.npz"), **out)
print("wrote", len(out), "arrays;", {k: float(v) for k, v in out.items() if k.endswith("_loss")})
