"""Golden vectors for the image-space losses either side of the rasterizer, made with the
reference's own `l1_loss` (src/entities/losses.py:8-27) combined exactly as the call sites combine it:
src/entities/tracker.py:112-128 (tracking colour / depth losses, alpha**3 weighting, masks, the
`sum / count(>0)` normalisation) and src/entities/mapper.py:160-165 (masked L1 means), with torch autograd
gradients.  Authoring container only:

    python tests/golden/make_golden_losses.py   ->  tests/golden/losses_reference.npz
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from src.entities.losses import l1_loss  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
H, W = 52, 70


def images(seed):
    g = torch.Generator().manual_seed(seed)
    color = torch.rand(3, H, W, generator=g)
    depth = 1 + 3 * torch.rand(1, H, W, generator=g)
    alpha = torch.rand(1, H, W, generator=g)
    gt_c = torch.rand(3, H, W, generator=g)
    gt_d = 1 + 3 * torch.rand(1, H, W, generator=g)
    gt_d[:, : H // 7] = 0                     # invalid depth rows
    color[:, 3, 5] = gt_c[:, 3, 5]            # exact zeros: the +1e-8 keeps them counted
    return color, depth, alpha, gt_c, gt_d


def main():
    out = {}
    color, depth, alpha, gt_c, gt_d = images(31)
    for k, v in (("color", color), ("depth", depth), ("alpha", alpha), ("gt_color", gt_c), ("gt_depth", gt_d)):
        out[k] = v.numpy()
    depth_mask = gt_d > 0
    tracking_mask = depth_mask & (alpha > 0.2)          # tracker.py:103-106 (filter_alpha)
    out["tracking_mask"] = tracking_mask.numpy()
    for tag, soft_alpha, mask_color in (("soft", True, False), ("soft_masked", True, True), ("hard", False, False)):
        c, d, a = (t.clone().requires_grad_(True) for t in (color, depth, alpha))
        # tracker.py:112-126
        color_loss = l1_loss(c, gt_c, agg="none") + 1e-8
        depth_loss = (l1_loss(d, gt_d, agg="none") + 1e-8) * tracking_mask
        if soft_alpha:
            a3 = a ** 3
            color_loss = color_loss * a3
            depth_loss = depth_loss * a3
            if mask_color:
                color_loss = color_loss * tracking_mask
        else:
            color_loss = color_loss * tracking_mask
        color_loss = color_loss.sum() / torch.sum(color_loss > 0)
        depth_loss = depth_loss.sum() / torch.sum(depth_loss > 0)
        (0.7 * color_loss + 1.3 * depth_loss).backward()
        out[f"track_{tag}_color_loss"] = color_loss.detach().numpy()
        out[f"track_{tag}_depth_loss"] = depth_loss.detach().numpy()
        out[f"track_{tag}_dcolor"] = c.grad.numpy()
        out[f"track_{tag}_ddepth"] = d.grad.numpy()
        out[f"track_{tag}_dalpha"] = a.grad.numpy() if a.grad is not None else np.zeros((1, H, W), np.float32)
    # mapper.py:160-165
    c, d = color.clone().requires_grad_(True), depth.clone().requires_grad_(True)
    mask = (gt_d > 0).squeeze(0) & (~torch.isnan(d)).squeeze(0)
    lc = l1_loss(c[:, mask], gt_c[:, mask])
    ld = l1_loss(d[:, mask], gt_d[0][mask])
    (lc + ld).backward()
    out["map_mask"] = mask.numpy()
    out["map_color_l1"], out["map_depth_l1"] = lc.detach().numpy(), ld.detach().numpy()
    out["map_dcolor"], out["map_ddepth"] = c.grad.numpy(), d.grad.numpy()
    path = os.path.join(HERE, "losses_reference.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
