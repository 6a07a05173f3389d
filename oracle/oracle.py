"""ctypes front-end of ``eags_oracle.c`` (the CPU restatement; TEST INFRASTRUCTURE ONLY).

``forward`` runs SURVEY Appendix A.1-A.3 and returns *every* intermediate the parity tests compare
(radii, tiles_touched, offsets, sorted keys/values, ranges, final_T, n_contrib, ...);
``backward`` runs A.4-A.5.  Arrays are numpy float32 / int32 / uint32 / uint64, C-contiguous.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "eags_oracle.c")
_SO = os.path.join(_HERE, "libeags_oracle.so")
_lib = None


class OracleSettings(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("M", C.c_int32), ("degree", C.c_int32),
        ("W", C.c_int32), ("H", C.c_int32), ("variant", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("view", C.c_float * 16), ("proj", C.c_float * 16),
        ("campos", C.c_float * 3), ("bg", C.c_float * 3),
    ]


def lib_path() -> str:
    return _SO


def build(force: bool = False) -> str:
    """Compile the C restatement with gcc (no FMA contraction, OpenMP)."""
    if (not force) and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(_SRC):
        return _SO
    cmd = ["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off",
           "-fno-fast-math", "-o", _SO, _SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.oracle_scan.restype = C.c_int64
        _lib.oracle_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(_load().oracle_num_threads())


def set_num_threads(n: int) -> None:
    _load().oracle_set_num_threads(C.c_int(int(n)))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def make_settings(P, M, degree, W, H, tanfovx, tanfovy, view, proj, campos, bg=(0, 0, 0),
                  scale_modifier=1.0, variant=0) -> OracleSettings:
    s = OracleSettings()
    s.P, s.M, s.degree, s.W, s.H, s.variant = int(P), int(M), int(degree), int(W), int(H), int(variant)
    s.tanfovx, s.tanfovy, s.scale_modifier = float(tanfovx), float(tanfovy), float(scale_modifier)
    s.view[:] = [float(v) for v in np.asarray(view, dtype=np.float32).reshape(-1)]
    s.proj[:] = [float(v) for v in np.asarray(proj, dtype=np.float32).reshape(-1)]
    s.campos[:] = [float(v) for v in np.asarray(campos, dtype=np.float32).reshape(-1)]
    s.bg[:] = [float(v) for v in np.asarray(bg, dtype=np.float32).reshape(-1)]
    return s


def forward(W, H, tanfovx, tanfovy, view, proj, campos, means3D, scales, rots, opacities,
            shs=None, colors_precomp=None, degree=0, bg=(0, 0, 0), scale_modifier=1.0,
            variant=0, want_n_touched=None, skip_blend=False):
    """A.1-A.3.  Returns a dict with all intermediates and outputs."""
    lib = _load()
    means3D, scales, rots = _f32(means3D), _f32(scales), _f32(rots)
    opacities = _f32(opacities).reshape(-1)
    shs, colors_precomp = _f32(shs), _f32(colors_precomp)
    P = means3D.shape[0]
    M = 1 if shs is None else shs.shape[1]
    if shs is None and colors_precomp is None:
        raise ValueError("need shs or colors_precomp")
    st = make_settings(P, M, degree, W, H, tanfovx, tanfovy, view, proj, campos, bg,
                       scale_modifier, variant)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    o = dict(settings=st, P=P, M=M, W=W, H=H, grid=(gx, gy))
    o["depths"] = np.zeros(P, np.float32)
    o["radii"] = np.zeros(P, np.int32)
    o["xy"] = np.zeros((P, 2), np.float32)
    o["conic_opacity"] = np.zeros((P, 4), np.float32)
    o["rgb"] = np.zeros((P, 3), np.float32)
    o["clamped"] = np.zeros((P, 3), np.uint8)
    o["rect"] = np.zeros((P, 4), np.int32)
    o["tiles_touched"] = np.zeros(P, np.uint32)
    o["cov3D"] = np.zeros((P, 6), np.float32)
    o["inputs"] = dict(means3D=means3D, scales=scales, rots=rots, opacities=opacities, shs=shs,
                       colors_precomp=colors_precomp)
    lib.oracle_preprocess(C.byref(st), _p(means3D), _p(scales), _p(rots), _p(opacities), _p(shs),
                          _p(colors_precomp), _p(o["depths"]), _p(o["radii"]), _p(o["xy"]),
                          _p(o["conic_opacity"]), _p(o["rgb"]), _p(o["clamped"]), _p(o["rect"]),
                          _p(o["tiles_touched"]), _p(o["cov3D"]))
    o["offsets"] = np.zeros(P, np.uint32)
    K = int(lib.oracle_scan(C.c_int32(P), _p(o["tiles_touched"]), _p(o["offsets"])))
    o["num_rendered"] = K
    o["keys"] = np.zeros(max(K, 1), np.uint64)[:K]
    o["point_list"] = np.zeros(max(K, 1), np.uint32)[:K]
    o["ranges"] = np.zeros((gx * gy, 2), np.uint32)
    keys_buf = np.zeros(max(K, 1), np.uint64)
    vals_buf = np.zeros(max(K, 1), np.uint32)
    rc = lib.oracle_binning(C.byref(st), _p(o["depths"]), _p(o["rect"]), _p(o["offsets"]),
                            C.c_int64(K), _p(keys_buf), _p(vals_buf), _p(o["ranges"]))
    if rc != 0:
        raise MemoryError("oracle_binning")
    o["keys"], o["point_list"] = keys_buf[:K], vals_buf[:K]
    if skip_blend:
        return o
    if want_n_touched is None:
        want_n_touched = variant == 1
    o["color"] = np.zeros((3, H, W), np.float32)
    o["depth"] = np.zeros((1, H, W), np.float32)
    o["third"] = np.zeros((1, H, W), np.float32)
    o["final_T"] = np.zeros((H, W), np.float32)
    o["n_contrib"] = np.zeros((H, W), np.uint32)
    o["fragile"] = np.zeros((H, W), np.uint8)
    o["n_touched"] = np.zeros(P, np.int32) if want_n_touched else None
    o["n_touched_fragile"] = np.zeros(P, np.uint8) if want_n_touched else None
    pairs = C.c_int64(0)
    lib.oracle_blend_forward(C.byref(st), _p(o["ranges"]), _p(vals_buf), _p(o["xy"]),
                             _p(o["conic_opacity"]), _p(o["rgb"]), _p(o["depths"]),
                             _p(o["color"]), _p(o["depth"]), _p(o["third"]), _p(o["final_T"]),
                             _p(o["n_contrib"]), _p(o["n_touched"]), _p(o["fragile"]),
                             _p(o["n_touched_fragile"]), C.byref(pairs))
    o["pair_count"] = int(pairs.value)
    o["_vals_buf"] = vals_buf
    return o


def backward(fwd, dL_dcolor, dL_ddepth, dL_dthird=None, want_tau=False):
    """A.4-A.5 on the state returned by :func:`forward`."""
    lib = _load()
    st = fwd["settings"]
    P, M, W, H = fwd["P"], fwd["M"], fwd["W"], fwd["H"]
    inp = fwd["inputs"]
    dL_dcolor = _f32(dL_dcolor).reshape(3, H, W)
    dL_ddepth = _f32(dL_ddepth).reshape(H, W)
    dL_dthird = None if dL_dthird is None else _f32(dL_dthird).reshape(H, W)
    g = {}
    g["dL_dmean2D"] = np.zeros((P, 2), np.float32)
    g["dL_dconic"] = np.zeros((P, 3), np.float32)
    g["dL_dopacity"] = np.zeros(P, np.float32)
    g["dL_dcolor"] = np.zeros((P, 3), np.float32)
    g["dL_ddepth"] = np.zeros(P, np.float32)
    rc = lib.oracle_blend_backward(C.byref(st), _p(fwd["ranges"]), _p(fwd["_vals_buf"]),
                                   _p(fwd["xy"]), _p(fwd["conic_opacity"]), _p(fwd["rgb"]),
                                   _p(fwd["depths"]), _p(fwd["final_T"]), _p(fwd["n_contrib"]),
                                   _p(dL_dcolor), _p(dL_ddepth), _p(dL_dthird),
                                   _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]),
                                   _p(g["dL_dcolor"]), _p(g["dL_ddepth"]))
    if rc != 0:
        raise MemoryError("oracle_blend_backward")
    g["dL_dmeans3D"] = np.zeros((P, 3), np.float32)
    g["dL_dscales"] = np.zeros((P, 3), np.float32)
    g["dL_drots"] = np.zeros((P, 4), np.float32)
    use_sh = inp["colors_precomp"] is None
    g["dL_dshs"] = np.zeros((P, M, 3), np.float32) if use_sh else None
    g["dL_dcolors_precomp"] = None if use_sh else np.zeros((P, 3), np.float32)
    g["dL_dtau"] = np.zeros(6, np.float32) if want_tau else None
    lib.oracle_preprocess_backward(C.byref(st), _p(inp["means3D"]), _p(inp["scales"]),
                                   _p(inp["rots"]), _p(inp["shs"]), _p(fwd["radii"]),
                                   _p(fwd["clamped"]), _p(g["dL_dmean2D"]), _p(g["dL_dconic"]),
                                   _p(g["dL_dcolor"]), _p(g["dL_ddepth"]), _p(g["dL_dmeans3D"]),
                                   _p(g["dL_dscales"]), _p(g["dL_drots"]), _p(g["dL_dshs"]),
                                   _p(g["dL_dcolors_precomp"]), _p(g["dL_dtau"]))
    return g
