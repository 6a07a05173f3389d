"""Shared helpers for the parity tests (oracle is the checker, never the thing under test on GPU)."""
import numpy as np
import torch

import oracle
from eags_raster import scene


def oracle_forward(cam, sc, degree=0, bg=(0, 0, 0), variant=0, colors_precomp=None, scale_modifier=1.0,
                   skip_blend=False):
    return oracle.forward(cam.W, cam.H, cam.tanfovx, cam.tanfovy, cam.viewmatrix.numpy(),
                          cam.projmatrix.numpy(), cam.campos.numpy(), sc["means3D"].cpu().numpy(),
                          sc["scales"].cpu().numpy(), sc["rotations"].cpu().numpy(),
                          sc["opacities"].cpu().numpy(),
                          shs=None if colors_precomp is not None else sc["shs"].cpu().numpy(),
                          colors_precomp=colors_precomp, degree=degree, bg=bg, variant=variant,
                          scale_modifier=scale_modifier, skip_blend=skip_blend)


def rel_err(a, b):
    """max |a-b| / max |b| (tensor-level relative error)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def assert_close_rel(a, b, rtol, atol_frac, what=""):
    """elementwise |a-b| <= rtol*|b| + atol_frac*max|b|"""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    tol = rtol * np.abs(b) + atol_frac * max(np.abs(b).max(), 1e-30)
    bad = np.abs(a - b) > tol
    assert not bad.any(), (f"{what}: {bad.sum()} / {bad.size} elements out of tolerance; worst abs err "
                           f"{np.abs(a - b).max():.3e} (max|ref| {np.abs(b).max():.3e})")


def assert_pixels_rel(a, b, rtol=1e-5, floor=1e-3, what=""):
    """TRUE elementwise relative error: |a-b| <= rtol * |b| wherever |b| > floor (no absolute slack
    borrowed from the brightest pixel), and |a-b| <= rtol * floor below it.  Returns the worst relative
    error seen above the floor (for the test log)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    big = np.abs(b) > floor
    err = np.abs(a - b)
    rel = np.where(big, err / np.maximum(np.abs(b), floor), 0.0)
    bad = (big & (rel > rtol)) | (~big & (err > rtol * floor))
    assert not bad.any(), (f"{what}: {bad.sum()} / {bad.size} elements exceed {rtol:g} relative "
                           f"(worst {rel.max():.3e} above the {floor:g} floor, worst abs below it "
                           f"{(err * ~big).max():.3e})")
    return float(rel.max())


def assert_rows_rel(a, b, rtol=1e-3, row_floor=1e-3, what=""):
    """Row-wise gradient check: every row (one Gaussian) whose reference max-norm exceeds
    ``row_floor * max|b|`` must agree to ``rtol`` relative to ITS OWN max-norm; smaller rows to
    ``rtol * row_floor * max|b|`` absolute.  Returns the worst per-row relative error."""
    a = np.asarray(a, dtype=np.float64).reshape(len(a), -1)
    b = np.asarray(b, dtype=np.float64).reshape(len(b), -1)
    gmax = max(np.abs(b).max(), 1e-30)
    rown = np.abs(b).max(axis=1)
    err = np.abs(a - b).max(axis=1)
    big = rown > row_floor * gmax
    rel = np.where(big, err / np.maximum(rown, 1e-30), 0.0)
    bad = (big & (rel > rtol)) | (~big & (err > rtol * row_floor * gmax))
    assert not bad.any(), (f"{what}: {bad.sum()} / {len(bad)} rows exceed {rtol:g} (worst row-relative "
                           f"{rel.max():.3e}; {big.sum()} rows above the floor)")
    return float(rel.max())


# ---- GPU-side helpers (the CUDA path under test, called through the C ABI binding) ----------------
class Settings:
    """Duck-typed GaussianRasterizationSettings built from a scene.Camera (device tensors)."""

    def __init__(self, cam, device, sh_degree=0, bg=(0, 0, 0), scale_modifier=1.0):
        self.image_height, self.image_width = cam.H, cam.W
        self.tanfovx, self.tanfovy = cam.tanfovx, cam.tanfovy
        self.bg = torch.tensor(bg, dtype=torch.float32, device=device)
        self.scale_modifier = scale_modifier
        self.viewmatrix = cam.viewmatrix.to(device)
        self.projmatrix = cam.projmatrix.to(device)
        self.projmatrix_raw = cam.projmatrix_raw.to(device)
        self.sh_degree = sh_degree
        self.campos = cam.campos.to(device)
        self.prefiltered = False
        self.debug = False


def cuda_forward(cam, sc, device="cuda", degree=0, bg=(0, 0, 0), variant=0, colors_precomp=None,
                 cov3D_precomp=None, scale_modifier=1.0, with_keys=True):
    from eags_raster import rasterize_forward
    st = Settings(cam, device, degree, bg, scale_modifier)
    d = {k: v.to(device) for k, v in sc.items()}
    out = rasterize_forward(
        st, variant, d["means3D"], d["opacities"],
        shs=None if colors_precomp is not None else d["shs"],
        colors_precomp=None if colors_precomp is None else torch.as_tensor(colors_precomp).to(device),
        scales=None if cov3D_precomp is not None else d["scales"],
        rotations=None if cov3D_precomp is not None else d["rotations"],
        cov3D_precomp=None if cov3D_precomp is None else torch.as_tensor(cov3D_precomp).to(device),
        want_n_touched=(variant == 1), with_keys=with_keys)
    return st, d, out


def nonfragile(o):
    return o["fragile"] == 0
