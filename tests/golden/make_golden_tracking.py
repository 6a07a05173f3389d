"""Golden vectors for the tracking pose transform made by the reference's OWN
functions: src/utils/tracker_utils.py (multiply_quaternions, compute_camera_opt_params,
transformation_to_quaternion) and src/utils/gaussian_model_utils.py (build_rotation), combined exactly as
src/entities/tracker.py:85-97 combines them.  Authoring container only (needs /root/reference):

    python tests/golden/make_golden_tracking.py   ->  tests/golden/tracking_reference.npz

`src.utils.utils` (open3d / rasterizer imports) is not importable here; tracker_utils needs only its
np2torch, which is stood in for below (utils.py: torch.from_numpy(array).float() moved to a device).
"""
import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference"
sys.path.insert(0, REF)
_stub = types.ModuleType("src.utils.utils")
_stub.np2torch = lambda array, device="cpu": torch.from_numpy(np.asarray(array)).float()
sys.modules["src.utils.utils"] = _stub
from src.utils.gaussian_model_utils import build_rotation  # noqa: E402
from src.utils.tracker_utils import (compute_camera_opt_params, multiply_quaternions,  # noqa: E402
                                     transformation_to_quaternion)

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    g = torch.Generator().manual_seed(77)
    out = {}
    n, cases = 48, 6
    pts = torch.randn(n, 3, generator=g) * 1.5 + torch.tensor([0.0, 0.0, 2.5])
    rots = F.normalize(torch.randn(n, 4, generator=g))
    out["pts"], out["rots"] = pts.numpy(), rots.numpy()
    # un-normalised optimisable quaternions (opt_cam_rot) and translations, small and large motions
    cam_rot = torch.randn(cases, 4, generator=g) * torch.tensor([0.01, 0.01, 0.3, 0.3, 1.0, 1.0])[:, None]
    cam_rot[:, 0] += torch.tensor([1.0, 1.7, 1.0, -0.8, 0.2, 0.0])
    cam_trans = torch.randn(cases, 3, generator=g) * torch.tensor([0.01, 0.01, 0.2, 0.2, 1.0, 1.0])[:, None]
    out["cam_rot"], out["cam_trans"] = cam_rot.numpy(), cam_trans.numpy()
    rel, moved, rq = [], [], []
    for c in range(cases):
        opt_cam_rot, opt_cam_trans = cam_rot[c], cam_trans[c]
        # tracker.py:85-97, verbatim in structure
        rel_transform = torch.eye(4, dtype=torch.float)
        rel_transform[:3, :3] = build_rotation(F.normalize(opt_cam_rot[None]), device="cpu")[0]
        rel_transform[:3, 3] = opt_cam_trans
        pts4 = torch.cat((pts, torch.ones(n, 1)), dim=1)
        transformed_pts = (rel_transform @ pts4.T).T[:, :3]
        quat = F.normalize(opt_cam_rot[None])
        _rotations = multiply_quaternions(rots, quat.unsqueeze(0)).squeeze(0)
        rel.append(rel_transform.numpy())
        moved.append(transformed_pts.numpy())
        rq.append(_rotations.numpy())
    out["rel_transform"], out["transformed_pts"], out["moved_rotations"] = np.stack(rel), np.stack(moved), np.stack(rq)
    # plain quaternion products
    qa, qb = torch.randn(32, 4, generator=g), torch.randn(32, 4, generator=g)
    out["qa"], out["qb"] = qa.numpy(), qb.numpy()
    out["qa_times_qb_ref"] = multiply_quaternions(qa, qb).numpy()      # multiply_quaternions(q, r)
    # compute_camera_opt_params / transformation_to_quaternion on rigid transforms
    w2c = np.stack(rel).astype(np.float64)
    out["opt_rot"] = np.stack([compute_camera_opt_params(m, device="cpu")[0].detach().numpy() for m in w2c])
    out["opt_trans"] = np.stack([compute_camera_opt_params(m, device="cpu")[1].detach().numpy() for m in w2c])
    out["rt_quat"] = np.stack([transformation_to_quaternion(m[:3]).numpy() for m in w2c])
    path = os.path.join(HERE, "tracking_reference.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", sorted(out))


if __name__ == "__main__":
    main()
