"""Loader of tests/golden/raster_oracle.npz (made by tests/golden/make_golden_raster.py)."""
import os
from types import SimpleNamespace

import numpy as np
import torch

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "raster_oracle.npz")
CASES = ("a", "b")


def load(name):
    z = np.load(PATH)
    g = {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(name + "/")}
    W, H, degree, variant, M, K = (int(v) for v in g["meta"])
    c = g["cam"]
    cam = SimpleNamespace(W=W, H=H, tanfovx=float(c[0]), tanfovy=float(c[1]),
                          viewmatrix=torch.from_numpy(c[2:18].reshape(4, 4).copy()),
                          projmatrix=torch.from_numpy(c[18:34].reshape(4, 4).copy()),
                          projmatrix_raw=torch.from_numpy(c[34:50].reshape(4, 4).copy()),
                          campos=torch.from_numpy(c[50:53].copy()))
    bg = tuple(float(v) for v in c[53:56])
    sc = {k: torch.from_numpy(g["in_" + k].copy()) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    return g, cam, sc, dict(degree=degree, variant=variant, M=M, K=K, bg=bg)
