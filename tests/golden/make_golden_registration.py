"""Generate tests/golden/registration_reference.npz from the reference's own registration loss and optimiser step.

Run in the authoring container only (needs /root/reference):  python tests/golden/make_golden_registration.py

Reference code executed (unchanged, on CPU):
  * ``get_loss_tracking_rgbd`` (src/gsr/loss.py:80-95, through ``get_loss_tracking_rgb`` :66-77) on seeded
    images, with its autograd gradients w.r.t. image / depth / opacity  -> pins eags_raster.losses.registration_loss
  * the optimiser side of ``viewpoint_localizer`` (src/gsr/solver.py:38-51,83-91): torch.optim.Adam on
    ``cam_rot_delta`` (lr 3e-3) / ``cam_trans_delta`` (lr 1e-3) with given gradients, then ``update_pose``
    (src/utils/pose_utils.py:76-93) on a ``Camera`` (src/gsr/camera.py), for 6 consecutive iterations
    -> pins eags_pose_adam_se3_update (eags_raster.tracking.PoseAdamSE3)
"""
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from src.gsr.camera import Camera  # noqa: E402
from src.gsr.loss import get_loss_tracking_rgbd  # noqa: E402
from src.utils.graphics_utils import getProjectionMatrix2  # noqa: E402
from src.utils.pose_utils import SE3_exp, update_pose  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    g = torch.Generator().manual_seed(5)
    out = {}
    # ---- the loss
    H, W = 52, 70
    for case in range(3):
        image = torch.rand(3, H, W, generator=g).requires_grad_(True)
        depth = (1.0 + 3.0 * torch.rand(1, H, W, generator=g)).requires_grad_(True)
        opacity = torch.rand(1, H, W, generator=g)
        opacity = torch.where(torch.rand(1, H, W, generator=g) < 0.5, 0.95 + 0.05 * opacity, opacity).requires_grad_(True)
        gt = torch.rand(3, H, W, generator=g)
        gt[:, torch.rand(H, W, generator=g) < 0.1] = 0.0                 # below rgb_boundary_threshold
        gt_depth = 1.0 + 3.0 * torch.rand(H, W, generator=g)
        gt_depth[torch.rand(H, W, generator=g) < 0.15] = 0.0             # invalid depth
        grad_mask = (torch.rand(1, H, W, generator=g) < 0.8) if case == 2 else torch.ones(1, H, W, dtype=torch.bool)
        vp = SimpleNamespace(original_image=gt, depth=gt_depth.numpy(), grad_mask=grad_mask)
        cfg = {"Training": {"monocular": False, "rgb_boundary_threshold": 0.01}}
        loss = get_loss_tracking_rgbd(cfg, image, depth, opacity, vp)
        loss.backward()
        for k, v in dict(image=image, depth=depth, opacity=opacity, gt=gt, gt_depth=gt_depth, grad_mask=grad_mask.float(),
                         loss=loss.detach().reshape(1), g_image=image.grad, g_depth=depth.grad,
                         g_opacity=opacity.grad).items():
            out[f"loss{case}_{k}"] = v.detach().numpy().astype(np.float32)
    # ---- Adam + update_pose
    fx, fy, cx, cy, Wc, Hc = 517.3, 516.5, 318.6, 255.3, 640, 480
    proj = getProjectionMatrix2(znear=0.01, zfar=100.0, fx=fx, fy=fy, cx=cx, cy=cy, W=Wc, H=Hc, device="cpu").transpose(0, 1)
    w2c = SE3_exp(torch.cat([torch.randn(3, generator=g), 0.3 * torch.randn(3, generator=g)]))
    cam = Camera(0, None, None, w2c, proj, fx, fy, cx, cy, 1.0, 1.0, Hc, Wc, device="cpu")
    cam.R, cam.T = w2c[:3, :3].clone(), w2c[:3, 3].clone()
    base_lr = 1e-3
    opt = torch.optim.Adam([{"params": [cam.cam_rot_delta], "lr": 3 * base_lr}, {"params": [cam.cam_trans_delta], "lr": base_lr}])
    grads, views, fulls, centers = [], [], [], []
    for it in range(6):
        gr = torch.randn(6, generator=g) * (10.0 ** float(torch.randint(-4, 0, (1,), generator=g)))
        opt.zero_grad()
        cam.cam_rot_delta.grad = gr[:3].clone()
        cam.cam_trans_delta.grad = gr[3:].clone()
        with torch.no_grad():
            opt.step()
            update_pose(cam)
        grads.append(gr)
        views.append(cam.world_view_transform.detach().clone())
        fulls.append(cam.full_proj_transform.detach().clone())
        centers.append(cam.camera_center.detach().clone())
    out.update(pose_w2c=w2c.numpy(), pose_proj_raw=proj.numpy(), pose_grads=torch.stack(grads).numpy(),
               pose_views=torch.stack(views).numpy(), pose_fulls=torch.stack(fulls).numpy(),
               pose_centers=torch.stack(centers).numpy())
    np.savez_compressed(os.path.join(HERE, "registration_reference.npz"), **{k: np.asarray(v, dtype=np.float32) for k, v in out.items()})
    print("wrote registration_reference.npz")


if __name__ == "__main__":
    main()
