"""Generate tests/golden/pose_update_reference.npz from the reference's own registration camera update.

Run in the authoring container only (needs /root/reference):  python tests/golden/make_golden_pose_update.py

Reference code executed (unchanged, on CPU): ``update_pose`` (src/utils/pose_utils.py:76-93) on a
``Camera`` (src/gsr/camera.py) whose ``cam_rot_delta`` / ``cam_trans_delta`` hold a given tau, then the
camera's ``world_view_transform`` / ``full_proj_transform`` / ``camera_center`` properties
(camera.py:97-111).  Cases: ordinary steps, steps below the 1e-5 small-angle switch, a chain of 20 steps.
Pins eags_pose_se3_update (csrc/pose.cu) -- tests/test_gpu_tracking.py.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from src.gsr.camera import Camera  # noqa: E402
from src.utils.graphics_utils import getProjectionMatrix2  # noqa: E402
from src.utils.pose_utils import SE3_exp, update_pose  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    g = torch.Generator().manual_seed(77)
    fx, fy, cx, cy, W, H = 517.3, 516.5, 318.6, 255.3, 640, 480
    proj = getProjectionMatrix2(znear=0.01, zfar=100.0, fx=fx, fy=fy, cx=cx, cy=cy, W=W, H=H, device="cpu").transpose(0, 1)
    cases = []
    for i in range(12):
        w2c = SE3_exp(torch.cat([torch.randn(3, generator=g), 0.5 * torch.randn(3, generator=g)]))
        scale = [1e-2, 1e-3, 1e-7, 0.3][i % 4]           # 1e-7: below the small-angle switch of SO3_exp / V
        taus = [scale * torch.randn(6, generator=g) for _ in range(20 if i == 11 else 1)]
        cam = Camera(i, None, None, w2c, proj, fx, fy, cx, cy, 1.0, 1.0, H, W, device="cpu")
        cam.R, cam.T = w2c[:3, :3].clone(), w2c[:3, 3].clone()
        for tau in taus:
            cam.cam_trans_delta.data.copy_(tau[:3])
            cam.cam_rot_delta.data.copy_(tau[3:])
            update_pose(cam)
        cases.append(dict(w2c=w2c.numpy(), taus=torch.stack(taus).numpy(),
                          view=cam.world_view_transform.detach().numpy(),
                          full=cam.full_proj_transform.detach().numpy(),
                          center=cam.camera_center.detach().numpy()))
    out = {"proj_raw": proj.numpy(), "n": np.array(len(cases))}
    for i, c in enumerate(cases):
        for k, v in c.items():
            out[f"{k}_{i}"] = v.astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "pose_update_reference.npz"), **out)
    print("wrote pose_update_reference.npz:", len(cases), "cases")


if __name__ == "__main__":
    main()
