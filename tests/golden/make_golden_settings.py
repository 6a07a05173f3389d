"""Golden vectors for the per-view settings made by the reference's OWN
`get_render_settings` (src/utils/utils.py:77-117), imported unchanged with THIS repo's `gaussian_rasterizer`
package standing where the un-vendored pip dependency would be - i.e. the drop-in import the reference performs
at src/utils/utils.py:8 resolves to the shim, and the returned object is the shim's NamedTuple.  open3d (not
installed here, unused by this function) is stood in for by an empty module.  Authoring container only:

    python tests/golden/make_golden_settings.py   ->  tests/golden/settings_reference.npz
"""
import math
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = ["/root/reference", os.path.join(ROOT, "eags-slam_b200"), ROOT]
_o3d = types.ModuleType("open3d")
_o3d.geometry = types.SimpleNamespace(PointCloud=object)     # only named in a type annotation (utils.py:49)
sys.modules.setdefault("open3d", _o3d)
import gaussian_rasterizer  # noqa: E402  (this repo's shim)
from src.utils.utils import get_render_settings  # noqa: E402
from src.gsr.camera import Camera  # noqa: E402
from src.utils.graphics_utils import focal2fov, getProjectionMatrix2  # noqa: E402
from eags_raster import scene  # noqa: E402

CAMS = [(640, 480, 517.306408, 516.469215, 318.643040, 255.313989),
        (1200, 680, 600.0, 600.0, 599.5, 339.5),
        (1752, 1168, 625.584, 626.960, 876.0, 584.0),
        (70, 52, 56.0, 55.0, 33.3, 27.1)]
ROTS = [(0, 0), (3, -2), (-4, 4), (25, 10)]


def main():
    out = {"cams": np.array(CAMS, np.float64), "rots": np.array(ROTS, np.float64)}
    for i, ((W, H, fx, fy, cx, cy), rot) in enumerate(zip(CAMS, ROTS)):
        w2c = scene.rot_xy(*rot).numpy().astype(np.float64)
        w2c[:3, 3] = [0.05 * i, -0.02 * i, 0.1 * i]
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float64)
        s = get_render_settings(W, H, K, w2c, device="cpu")
        assert type(s) is gaussian_rasterizer.GaussianRasterizationSettings
        out[f"{i}/w2c"] = w2c
        out[f"{i}/scalars"] = np.array([s.image_height, s.image_width, s.tanfovx, s.tanfovy, s.scale_modifier,
                                        s.sh_degree, float(s.prefiltered), float(s.debug)], np.float64)
        out[f"{i}/viewmatrix"] = s.viewmatrix.contiguous().numpy()
        out[f"{i}/projmatrix"] = s.projmatrix.contiguous().numpy()
        out[f"{i}/campos"] = s.campos.numpy()
        out[f"{i}/bg"] = s.bg.numpy()
    # variant B (loop-closure registration): src/gsr/camera.py properties as src/gsr/renderer.py:59-73 reads them,
    # projection matrix as src/entities/lc.py:70-77 builds it (fy := fx there; the fov keeps the real fy)
    for i, ((W, H, fx, fy, cx, cy), rot) in enumerate(zip(CAMS, ROTS)):
        W, H = int(W), int(H)
        proj = getProjectionMatrix2(znear=0.01, zfar=100.0, fx=fx, fy=fx, cx=cx, cy=cy, W=W, H=H, device="cpu").T
        fovx, fovy = focal2fov(fx, W), focal2fov(fy, H)
        cam = Camera(i, None, None, torch.eye(4), proj, fx, fy, cx, cy, fovx, fovy, H, W, device="cpu")
        w2c = torch.from_numpy(out[f"{i}/w2c"]).float()
        cam.update_RT(w2c[:3, :3], w2c[:3, 3])
        out[f"{i}/b_tanfov"] = np.array([math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)], np.float64)
        out[f"{i}/b_viewmatrix"] = cam.world_view_transform.contiguous().numpy()
        out[f"{i}/b_projmatrix"] = cam.full_proj_transform.contiguous().numpy()
        out[f"{i}/b_projmatrix_raw"] = cam.projection_matrix.contiguous().numpy()
        out[f"{i}/b_campos"] = cam.camera_center.numpy()
    path = os.path.join(HERE, "settings_reference.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
