"""Drop-in check against the reference checkout itself (authoring container only: skipped where
/root/reference does not exist, e.g. on the GPU box).  The reference's UNCHANGED src/utils/utils.py is imported
with this repo's packages answering `from gaussian_rasterizer import ...` (utils.py:8), its
get_render_settings builds the shim's settings tuple, and its render_gaussian_model (utils.py:120-186) reaches
the operator with the reference's own keyword set; likewise src/gsr/renderer.py::render (variant B, with
src/gsr/camera.py's Camera and `from simple_knn._C import distCUDA2` of gaussian_model.py:17).  Last, the submap
hand-off layout: the reference's own Submap.load reads what handoff.to_ckpt wrote and
handoff packs what the reference's Submap.save wrote.  There is no GPU here, so the call must fail LOUDLY in the
shim (no CPU fallback) - which also proves the reference's caller got all the way to the boundary.
Runs in a subprocess so the reference's `src` package and the open3d stand-in never enter the test process."""
import os
import subprocess
import sys
import textwrap

import pytest

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = textwrap.dedent("""
    import sys, types
    import numpy as np, torch
    o3d = types.ModuleType("open3d")
    o3d.geometry = types.SimpleNamespace(PointCloud=object)     # named in one type annotation only (utils.py:49)
    sys.modules["open3d"] = o3d
    sys.path[:0] = [%(ref)r, %(pkg)r, %(root)r]
    import gaussian_rasterizer as shim
    from src.utils import utils as ref_utils                    # the reference's own module, unchanged
    assert ref_utils.GaussianRasterizer is shim.GaussianRasterizer
    K = np.array([[517.3, 0, 318.6], [0, 516.5, 255.3], [0, 0, 1.0]])
    settings = ref_utils.get_render_settings(640, 480, K, np.eye(4), device="cpu")
    assert type(settings) is shim.GaussianRasterizationSettings and settings.image_height == 480

    class Model:                                                # what render_gaussian_model asks of a GaussianModel
        device = "cpu"
        def __init__(self, P):
            g = torch.Generator().manual_seed(0)
            self.xyz = torch.randn(P, 3, generator=g) + torch.tensor([0.0, 0.0, 3.0])
            self.sh = torch.rand(P, 1, 3, generator=g)
        def get_xyz(self): return self.xyz
        def get_opacity(self): return torch.full((self.xyz.shape[0], 1), 0.5)
        def get_features(self): return self.sh
        def get_scaling(self): return torch.full((self.xyz.shape[0], 3), 0.02)
        def get_rotation(self): return torch.tensor([[1.0, 0, 0, 0]]).repeat(self.xyz.shape[0], 1)

    try:
        ref_utils.render_gaussian_model(Model(100), settings)
    except RuntimeError as e:
        assert "CUDA" in str(e) and "no CPU" in str(e), str(e)
        print("LOUD:", e)
    else:
        raise SystemExit("render_gaussian_model returned on CPU tensors: a CPU fallback exists")

    # ---- variant B: src/gsr/renderer.py::render with its own Camera (loop-closure registration) ----
    ply = types.ModuleType("plyfile")
    ply.PlyData = ply.PlyElement = object                       # gaussian_model.py:16, unused on this path
    sys.modules["plyfile"] = ply
    import diff_gaussian_rasterization as shim_b
    import simple_knn._C as knn_shim
    from src.gsr import renderer as ref_renderer                # imports src.entities.gaussian_model -> simple_knn._C
    from src.entities import gaussian_model as ref_gm
    from src.gsr.camera import Camera
    from src.utils.graphics_utils import focal2fov, getProjectionMatrix2
    assert ref_renderer.GaussianRasterizer is shim_b.GaussianRasterizer
    assert ref_gm.distCUDA2 is knn_shim.distCUDA2
    proj = getProjectionMatrix2(znear=0.01, zfar=100.0, fx=517.3, fy=517.3, cx=318.6, cy=255.3, W=640, H=480,
                                device="cpu").T                 # lc.py:70-75
    cam = Camera(0, None, None, torch.eye(4), proj, 517.3, 516.5, 318.6, 255.3, focal2fov(517.3, 640),
                 focal2fov(516.5, 480), 480, 640, device="cpu")

    class ModelB(Model):
        active_sh_degree, max_sh_degree = 0, 0

    pipe = types.SimpleNamespace(compute_cov3D_python=False, convert_SHs_python=False)
    try:
        ref_renderer.render(cam, ModelB(100), pipe, torch.zeros(3), device="cpu")
    except RuntimeError as e:
        assert "CUDA" in str(e) and "no CPU" in str(e), str(e)
        print("LOUD-B:", e)
    else:
        raise SystemExit("gsr.renderer.render returned on CPU tensors: a CPU fallback exists")

    # ---- submap hand-off (SURVEY 8(f) N-4): the reference's own Submap.save / Submap.load on the same layout ----
    import os, pathlib, tempfile
    for name in ("roma", "wandb"):                              # imported by submap.py / io_utils.py, unused here
        sys.modules[name] = types.ModuleType(name)
    from src.entities.submap import Submap
    from eags_raster import handoff
    P = 321
    g = torch.Generator().manual_seed(9)
    sub = {"id": 7,
           "gaussian_params": {"xyz": torch.randn(P, 3, generator=g), "features_dc": torch.randn(P, 1, 3, generator=g),
                               "features_rest": torch.zeros(P, 0, 3), "scaling": torch.randn(P, 3, generator=g),
                               "rotation": torch.randn(P, 4, generator=g), "opacity": torch.randn(P, 1, generator=g)},
           "Tmc": torch.eye(4, dtype=torch.float64)[None].repeat(4, 1, 1), "kf_id": torch.arange(20, 24),
           "T_prev_m": torch.eye(4, dtype=torch.float64)}
    with tempfile.TemporaryDirectory() as d:
        path = handoff.to_ckpt(sub, d)                          # this repo writes ...
        back = Submap().load(path, device="cpu")                # ... the reference's loader reads (submap.py:21-37)
        assert back.id == 7 and torch.equal(back.kf_id, sub["kf_id"]) and torch.equal(back.Tmc, sub["Tmc"])
        assert all(torch.equal(back.gaussian_params[k], v) for k, v in sub["gaussian_params"].items())
        back.id = 8
        back.save(pathlib.Path(d) / "ref")                      # the reference writes (submap.py:109-124) ...
        theirs = torch.load(os.path.join(d, "ref", "000008.ckpt"), weights_only=False)
        assert tuple(theirs) == handoff.CKPT_KEYS
        handoff.check_layout(theirs)                            # ... this repo packs / unpacks
        again = handoff.unpack(*handoff.pack(theirs))
        assert again["id"] == 8 and all(torch.equal(again["gaussian_params"][k], v) for k, v in sub["gaussian_params"].items())
    print("HANDOFF ok")
""")


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference checkout not present")
def test_reference_callers_reach_the_shims_fail_loudly_without_a_gpu_and_share_the_submap_layout():
    code = SCRIPT % dict(ref=REF, pkg=os.path.join(ROOT, "eags-slam_b200"), root=ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "LOUD:" in r.stdout and "LOUD-B:" in r.stdout and "HANDOFF ok" in r.stdout
