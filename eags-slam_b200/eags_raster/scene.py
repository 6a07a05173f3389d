"""Seeded synthetic scenes + cameras for the benchmark workloads.

Camera conventions follow the reference glue: ``viewmatrix = w2c^T`` and
``projmatrix = viewmatrix @ opengl_proj^T`` (reference src/utils/utils.py:93-104), i.e. the same
matrices ``getProjectionMatrix2`` builds (reference src/utils/graphics_utils.py:72-93).
Everything here is host-side input generation; no rasterizer math.
"""
from __future__ import annotations

import math
from typing import Dict, NamedTuple

import torch

# name -> (W, H, fx, fy, cx, cy, default P)
WORKLOADS = {
    # reference configs/TUM_RGBD/rgbd_dataset_freiburg1_desk.yaml:8-13
    "c1": (640, 480, 517.306408, 516.469215, 318.643040, 255.313989, 50_000),
    "c2": (640, 480, 517.306408, 516.469215, 318.643040, 255.313989, 300_000),
    # reference configs/Replica/replica.yaml:49-55
    "c3": (1200, 680, 600.0, 600.0, 599.5, 339.5, 500_000),
    # 2x the 876x584 intrinsics of reference configs/scannetpp/281bc17764.yaml:9-14
    "c4": (1752, 1168, 625.584, 626.960, 876.0, 584.0, 1_500_000),
    "c5": (640, 480, 517.306408, 516.469215, 318.643040, 255.313989, 300_000),
    # the shapes the reference's variant-A call sites really pass (dataset size AFTER the
    # crop of src/entities/datasets.py:34-39, principal point shifted by the crop)
    "tum_crop": (540, 380, 517.306408, 516.469215, 318.643040 - 50, 255.313989 - 50, 200_000),     # crop_edge 50
    "scannet_crop": (616, 456, 577.590698, 578.729797, 318.905426 - 12, 242.683609 - 12, 200_000),  # scannet.yaml:49-56
    "scannetpp": (876, 584, 312.79197434640764, 313.48022477591036, 438.0, 292.0, 300_000),        # 281bc17764.yaml:9-14
    "replica": (1200, 680, 600.0, 600.0, 599.5, 339.5, 300_000),                                    # replica.yaml:49-55
    # test-only shapes (ragged: 70x52 is not a multiple of the 16x16 tile)
    "tiny": (64, 48, 51.7306408, 51.6469215, 31.8643040, 25.5313989, 300),
    "ragged": (70, 52, 56.0, 55.0, 33.3, 27.1, 400),
}


class Camera(NamedTuple):
    W: int
    H: int
    fx: float
    fy: float
    cx: float
    cy: float
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor      # (4,4) = w2c^T
    projmatrix: torch.Tensor      # (4,4) = viewmatrix @ P^T
    projmatrix_raw: torch.Tensor  # (4,4) = P^T
    campos: torch.Tensor          # (3,)
    w2c: torch.Tensor             # (4,4)


def make_camera(W, H, fx, fy, cx, cy, w2c=None, near=0.01, far=100.0) -> Camera:
    if w2c is None:
        w2c = torch.eye(4, dtype=torch.float32)
    w2c = w2c.to(torch.float32).cpu()
    P = torch.tensor(
        [[2 * fx / W, 0.0, -(W - 2 * cx) / W, 0.0],
         [0.0, 2 * fy / H, -(H - 2 * cy) / H, 0.0],
         [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
         [0.0, 0.0, 1.0, 0.0]], dtype=torch.float32)
    view = w2c.t().contiguous()
    proj_raw = P.t().contiguous()
    full = (view @ proj_raw).contiguous()
    campos = torch.linalg.inv(w2c)[:3, 3].contiguous()
    return Camera(W, H, fx, fy, cx, cy, W / (2 * fx), H / (2 * fy), view, full, proj_raw, campos, w2c)


def rot_xy(deg_x: float, deg_y: float) -> torch.Tensor:
    """w2c of a camera at the origin rotated by the given angles about x then y."""
    ax, ay = math.radians(deg_x), math.radians(deg_y)
    Rx = torch.tensor([[1, 0, 0], [0, math.cos(ax), -math.sin(ax)], [0, math.sin(ax), math.cos(ax)]])
    Ry = torch.tensor([[math.cos(ay), 0, math.sin(ay)], [0, 1, 0], [-math.sin(ay), 0, math.cos(ay)]])
    T = torch.eye(4)
    T[:3, :3] = (Ry @ Rx).to(torch.float32)
    return T


def keyframe_views(workload: str, n_views: int = 8):
    """The c4 keyframe batch: identity camera rotated +-4 degrees about x / y."""
    W, H, fx, fy, cx, cy, _ = WORKLOADS[workload]
    angles = [(0, 0), (4, 0), (-4, 0), (0, 4), (0, -4), (4, 4), (-4, -4), (4, -4)]
    return [make_camera(W, H, fx, fy, cx, cy, rot_xy(*angles[i % len(angles)])) for i in range(n_views)]


def make_scene(workload: str = "c2", P: int | None = None, seed: int = 0, M: int = 1,
               device="cpu", outside: float = 0.05, behind: float = 0.02,
               scale_range=(0.7, 2.5), order: str = "random") -> Dict[str, torch.Tensor]:
    """Synthetic Gaussians per SURVEY 8(d): pixels uniform in the image (+5 % outside up to
    1.5x the frustum, 2 % behind the camera), depth U(1,4) m, per-axis scale (z/fx)*U(0.7,2.5),
    random unit quaternions, opacity sigmoid(N(0,1.5^2)), SH-DC U(-1.7,1.7).

    ``order`` is the storage order of the same Gaussians: "random" (neighbours in memory
    land anywhere in the image -- the adversarial case for tile binning), "raster" (by seeding pixel,
    row-major: how the reference seeds a keyframe's Gaussians from its pixels, mapper.py:225-260 /
    gaussian_model.py:586-592) or "tile" (by 16x16 tile of the seeding pixel, see
    :func:`tile_order`)."""
    W, H, fx, fy, cx, cy, P0 = WORKLOADS[workload]
    P = int(P0 if P is None else P)
    g = torch.Generator().manual_seed(int(seed))
    u = torch.rand(P, generator=g) * W
    v = torch.rand(P, generator=g) * H
    z = 1.0 + 3.0 * torch.rand(P, generator=g)
    sel = torch.rand(P, generator=g)
    f_out, f_beh = float(outside), float(behind)
    outside = sel < f_out
    behind = (sel >= f_out) & (sel < f_out + f_beh)
    uo = (torch.rand(P, generator=g) * 1.5 - 0.25) * W
    vo = (torch.rand(P, generator=g) * 1.5 - 0.25) * H
    u = torch.where(outside, uo, u)
    v = torch.where(outside, vo, v)
    z = torch.where(behind, -z, z)
    x = (u + 0.5 - cx) / fx * z
    y = (v + 0.5 - cy) / fy * z
    means3D = torch.stack([x, y, z], dim=1)
    scales = (z.abs() / fx)[:, None] * (scale_range[0] + (scale_range[1] - scale_range[0]) * torch.rand(P, 3, generator=g))
    q = torch.randn(P, 4, generator=g)
    rotations = q / q.norm(dim=1, keepdim=True)
    opacities = torch.sigmoid(1.5 * torch.randn(P, 1, generator=g))
    shs = torch.zeros(P, M, 3)
    shs[:, 0, :] = torch.rand(P, 3, generator=g) * 3.4 - 1.7
    if M > 1:
        shs[:, 1:, :] = 0.1 * torch.randn(P, M - 1, 3, generator=g)
    gt_color = torch.rand(3, H, W, generator=g)
    gt_depth = 1.0 + 3.0 * torch.rand(1, H, W, generator=g)
    out = dict(means3D=means3D, scales=scales, rotations=rotations, opacities=opacities, shs=shs)
    if order != "random":
        ui, vi = u.clamp(0, W - 1).long(), v.clamp(0, H - 1).long()
        if order == "raster":
            key = vi * W + ui
        elif order == "tile":
            key = ((vi // 16) * ((W + 15) // 16) + ui // 16) * 256 + (vi % 16) * 16 + ui % 16
        else:
            raise ValueError("order must be 'random', 'raster' or 'tile'")
        perm = torch.argsort(key, stable=True)
        out = {k: t[perm] for k, t in out.items()}
    out.update(gt_color=gt_color, gt_depth=gt_depth)
    return {k: t.to(torch.float32).contiguous().to(device) for k, t in out.items()}


def tile_order(means3D: torch.Tensor, cam: "Camera") -> torch.Tensor:
    """Permutation that stores a submap's Gaussians by the 16x16 tile their mean projects to in ``cam``
    (Gaussians outside the image or behind the camera last).  Applying it to every per-Gaussian tensor
    (and the optimiser state) changes no rendered value -- the rasterizer's output order is (depth, index)
    per tile and ties are measure-zero -- but lets preprocess_fwd's CTA-level aggregation reserve whole
    runs of a tile's bucket with one atomic (csrc/preprocess.cu)."""
    dev = means3D.device
    w2c = cam.w2c.to(dev)
    pc = means3D @ w2c[:3, :3].T + w2c[:3, 3]
    z = pc[:, 2].clamp_min(1e-6)
    ui = (pc[:, 0] / z * cam.fx + cam.cx - 0.5).floor().long()
    vi = (pc[:, 1] / z * cam.fy + cam.cy - 0.5).floor().long()
    inside = (pc[:, 2] > 0.2) & (ui >= 0) & (ui < cam.W) & (vi >= 0) & (vi < cam.H)
    gx = (cam.W + 15) // 16
    key = torch.where(inside, (vi // 16) * gx + ui // 16, torch.full_like(ui, 1 << 40))
    return torch.argsort(key, stable=True)


def workload_camera(workload: str, w2c=None) -> Camera:
    W, H, fx, fy, cx, cy, _ = WORKLOADS[workload]
    return make_camera(W, H, fx, fy, cx, cy, w2c)
