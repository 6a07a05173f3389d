"""Frozen input/output vectors of the rasterizer path on two small seeded scenes (variant A and B).

The reference's own kernels are un-vendored, so these vectors come from the CPU
oracle (oracle/eags_oracle.c) and are NOT a pin by the reference: they freeze the checker (compiler /
flag drift shows up in tests/test_oracle_golden_raster.py) and give the CUDA path a fixed target that
does not depend on the oracle being built on the GPU box (tests/test_gpu_golden.py).

    python tests/golden/make_golden_raster.py      ->  tests/golden/raster_oracle.npz
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import oracle  # noqa: E402
from eags_raster import scene  # noqa: E402

CASES = {
    # name: (workload, P, M, degree, variant, bg, camera rotation (deg about x, y), seed)
    "a": ("ragged", 700, 1, 0, 0, (0.0, 0.0, 0.0), (0, 0), 21),
    "b": ("ragged", 500, 4, 1, 1, (0.1, 0.2, 0.3), (3, -2), 22),
}
FWD = ("radii", "tiles_touched", "ranges", "point_list", "keys", "xy", "depths", "rgb", "conic_opacity", "clamped",
       "color", "depth", "third", "final_T", "n_contrib", "fragile")
BWD = ("dL_dmean2D", "dL_dmeans3D", "dL_dscales", "dL_drots", "dL_dopacity", "dL_dshs", "dL_dtau")


def main():
    out = {}
    for name, (wl, P, M, degree, variant, bg, rot, seed) in CASES.items():
        sc = scene.make_scene(wl, P=P, seed=seed, M=M)
        cam = scene.workload_camera(wl, w2c=scene.rot_xy(*rot))
        o = oracle.forward(cam.W, cam.H, cam.tanfovx, cam.tanfovy, cam.viewmatrix.numpy(), cam.projmatrix.numpy(),
                           cam.campos.numpy(), sc["means3D"].numpy(), sc["scales"].numpy(), sc["rotations"].numpy(),
                           sc["opacities"].numpy(), shs=sc["shs"].numpy(), degree=degree, bg=bg, variant=variant)
        g = torch.Generator().manual_seed(seed + 100)
        dC, dD, dA = (torch.randn(c, cam.H, cam.W, generator=g).numpy() for c in (3, 1, 1))
        fr = o["fragile"] != 0
        dC[:, fr] = 0
        dD[:, fr] = 0
        dA[:, fr] = 0
        b = oracle.backward(o, dC, dD, dA, want_tau=True)
        pre = name + "/"
        out[pre + "meta"] = np.array([cam.W, cam.H, degree, variant, M, o["num_rendered"]], np.int64)
        out[pre + "cam"] = np.concatenate([np.array([cam.tanfovx, cam.tanfovy], np.float32),
                                           cam.viewmatrix.numpy().reshape(-1), cam.projmatrix.numpy().reshape(-1),
                                           cam.projmatrix_raw.numpy().reshape(-1), cam.campos.numpy().reshape(-1),
                                           np.array(bg, np.float32)]).astype(np.float32)
        for k in ("means3D", "scales", "rotations", "opacities", "shs"):
            out[pre + "in_" + k] = sc[k].numpy()
        out[pre + "dL_dcolor"], out[pre + "dL_ddepth"], out[pre + "dL_dthird"] = dC, dD, dA
        for k in FWD:
            out[pre + k] = o[k]
        if variant == 1:
            out[pre + "n_touched"] = o["n_touched"]
            out[pre + "n_touched_fragile"] = o["n_touched_fragile"]
        for k in BWD:
            out[pre + k] = b[k]
    path = os.path.join(HERE, "raster_oracle.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
