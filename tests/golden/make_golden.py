"""Generate tests/golden/reference_helpers.npz from the reference's own importable Python helpers.

Run in the authoring container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

The rasterizer kernels themselves are un-vendored pip dependencies, so the
only reference code that can execute here are the convention-defining helpers:
  src/utils/gaussian_model_utils.py : eval_sh, build_rotation, build_scaling_rotation, strip_symmetric
  src/utils/graphics_utils.py       : getProjectionMatrix2, getWorld2View2, focal2fov
  src/utils/pose_utils.py           : SE3_exp
Their outputs on seeded inputs are the golden vectors that pin the oracle's conventions
(tests/test_oracle_golden.py).
"""
import math
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
from src.utils.gaussian_model_utils import (build_rotation, build_scaling_rotation, eval_sh,  # noqa: E402
                                            strip_symmetric)
from src.utils.graphics_utils import focal2fov, getProjectionMatrix2, getWorld2View2  # noqa: E402
from src.utils.pose_utils import SE3_exp  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    g = torch.Generator().manual_seed(1234)
    out = {}
    n = 64
    q = torch.randn(n, 4, generator=g)
    q = q / q.norm(dim=1, keepdim=True)
    s = torch.rand(n, 3, generator=g) * 0.05 + 0.001
    out["quat"] = q.numpy()
    out["scale"] = s.numpy()
    out["R"] = build_rotation(q, device="cpu").numpy()
    # gaussian_model.py:70-74 build_covariance_from_scaling_rotation
    L = build_scaling_rotation(1.0 * s, q, "cpu")
    out["cov6"] = strip_symmetric(L @ L.transpose(1, 2), device="cpu").numpy()
    # SH: reference layout for eval_sh is [..., C, coeffs]
    sh = torch.randn(n, 3, 16, generator=g)
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=1, keepdim=True)
    out["sh_pc"] = sh.numpy()          # (n, 3, 16)
    out["dirs"] = d.numpy()
    for deg in range(4):
        rgb = eval_sh(deg, sh, d)
        out[f"sh_rgb_deg{deg}"] = torch.clamp_min(rgb + 0.5, 0.0).numpy()   # renderer.py:109-110
        out[f"sh_raw_deg{deg}"] = rgb.numpy()
    # projection matrices for the benchmark intrinsics (W,H,fx,fy,cx,cy)
    cams = np.array([[640, 480, 517.306408, 516.469215, 318.643040, 255.313989],
                     [1200, 680, 600.0, 600.0, 599.5, 339.5],
                     [1752, 1168, 625.584, 626.960, 876.0, 584.0],
                     [70, 52, 56.0, 55.0, 33.3, 27.1]], dtype=np.float64)
    out["cams"] = cams
    Ps, fovs = [], []
    for W, H, fx, fy, cx, cy in cams:
        P = getProjectionMatrix2(znear=0.01, zfar=100.0, cx=cx, cy=cy, fx=fx, fy=fy, W=W, H=H, device="cpu")
        Ps.append(P.numpy())
        fovs.append([math.tan(focal2fov(fx, W) / 2), math.tan(focal2fov(fy, H) / 2)])
    out["proj"] = np.stack(Ps)
    out["tanfov"] = np.array(fovs)
    # world->view + camera centre (camera.py:97-111)
    taus = torch.randn(8, 6, generator=g) * 0.2
    taus[0] *= 1e-7
    taus[1] = 0
    out["taus"] = taus.numpy()
    Ts = torch.stack([SE3_exp(t) for t in taus])
    out["se3"] = Ts.numpy()
    w2v = torch.stack([getWorld2View2(T[:3, :3], T[:3, 3]) for T in Ts])
    out["world2view2"] = w2v.numpy()
    np.savez_compressed(os.path.join(HERE, "reference_helpers.npz"), **out)
    print("wrote reference_helpers.npz with", sorted(out))


if __name__ == "__main__":
    main()
