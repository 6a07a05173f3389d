"""ctypes binding of ``libeags_raster.so`` (the C ABI of ``include/eags_raster.h``).

This is the whole "extension": plain pointers, sizes and a stream handle cross the boundary; PyTorch
only provides device memory and streams.  There is no CPU fallback -- if the library has not been
built (``python __graft_entry__.py`` / ``make -C eags-slam_b200/csrc``) importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# EAGS_RASTER_LIB selects another build of the same library (A/B measurements of kernel variants)
LIB_PATH = os.environ.get("EAGS_RASTER_LIB") or os.path.join(_HERE, "libeags_raster.so")

VARIANT_A = 0
VARIANT_B = 1

c_float_p = C.c_void_p  # device pointers travel as integers


class EagsView(C.Structure):
    _fields_ = [
        ("image_height", C.c_int32), ("image_width", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("sh_degree", C.c_int32), ("variant", C.c_int32), ("prefiltered", C.c_int32),
        ("debug", C.c_int32),
        ("bg", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("projmatrix_raw", C.c_void_p), ("campos", C.c_void_p),
    ]


class EagsGaussians(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("M", C.c_int32),
        ("means3D", C.c_void_p), ("opacities", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
        ("cov3D_precomp", C.c_void_p),
    ]


class EagsImages(C.Structure):
    _fields_ = [("color", C.c_void_p), ("depth", C.c_void_p), ("third", C.c_void_p),
                ("n_touched", C.c_void_p)]


class EagsImageGrads(C.Structure):
    _fields_ = [("dL_dcolor", C.c_void_p), ("dL_ddepth", C.c_void_p), ("dL_dthird", C.c_void_p)]


class EagsInputGrads(C.Structure):
    _fields_ = [
        ("dL_dmeans3D", C.c_void_p), ("dL_dmeans2D", C.c_void_p), ("dL_dopacities", C.c_void_p),
        ("dL_dshs", C.c_void_p), ("dL_dcolors_precomp", C.c_void_p), ("dL_dscales", C.c_void_p),
        ("dL_drotations", C.c_void_p), ("dL_dcov3D_precomp", C.c_void_p), ("dL_dtau", C.c_void_p),
    ]


class EagsImageL1(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("soft_alpha", C.c_int32), ("mask_color", C.c_int32),
        ("norm_mode", C.c_int32), ("eps", C.c_float),
        ("color", C.c_void_p), ("depth", C.c_void_p), ("alpha", C.c_void_p), ("gt_color", C.c_void_p),
        ("gt_depth", C.c_void_p), ("mask", C.c_void_p), ("mask_depth", C.c_void_p), ("alpha_min", C.c_float),
    ]


class EagsMappingLoss(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("lambda_dssim", C.c_float),
        ("image", C.c_void_p), ("depth", C.c_void_p), ("gt_image", C.c_void_p), ("gt_depth", C.c_void_p),
        ("mask", C.c_void_p), ("valid_depth_mask", C.c_int32),
    ]


# name -> (restype, argtypes); also the list tests/test_abi.py checks against include/eags_raster.h
SIGNATURES = {
    "eags_abi_version": (C.c_int, []),
    "eags_last_error": (C.c_char_p, []),
    "eags_launch_count": (C.c_uint64, []),
    "eags_geom_state_bytes": (C.c_size_t, [C.c_int32]),
    "eags_image_state_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "eags_binning_state_bytes": (C.c_size_t, [C.c_int64, C.c_int32]),
    "eags_backward_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "eags_backward_stats_offset": (C.c_size_t, [C.c_int32]),
    "eags_set_option": (C.c_int, [C.c_char_p, C.c_int32]),
    "eags_get_option": (C.c_int, [C.c_char_p, C.POINTER(C.c_int32)]),
    "eags_kernel_times": (C.c_int, [C.c_int32, C.POINTER(C.c_char_p), C.POINTER(C.c_float)]),
    "eags_forward_geom": (C.c_int, [C.POINTER(EagsView), C.POINTER(EagsGaussians), C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "eags_forward_render": (C.c_int, [C.POINTER(EagsView), C.POINTER(EagsGaussians), C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                      C.POINTER(EagsImages), C.c_void_p]),
    "eags_binning_state_bytes_bucketed": (C.c_size_t, [C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "eags_forward_geom_bucketed": (C.c_int, [C.POINTER(EagsView), C.POINTER(EagsGaussians), C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                             C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "eags_forward_render_bucketed": (C.c_int, [C.POINTER(EagsView), C.POINTER(EagsGaussians), C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                               C.c_int32, C.POINTER(EagsImages), C.c_void_p]),
    "eags_backward": (C.c_int, [C.POINTER(EagsView), C.POINTER(EagsGaussians), C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                C.POINTER(EagsImageGrads), C.POINTER(EagsInputGrads), C.c_void_p,
                                C.c_void_p]),
    "eags_peer_flags_bytes": (C.c_size_t, []),
    "eags_backward_routed": (C.c_int, [C.POINTER(EagsView), C.POINTER(EagsGaussians), C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                       C.POINTER(EagsImageGrads), C.POINTER(EagsInputGrads), C.c_void_p,
                                       C.c_void_p, C.c_void_p]),
    "eags_peer_reduce": (C.c_int, [C.c_void_p, C.c_uint32, C.c_float, C.c_int32, C.c_void_p]),
    "eags_peer_adam_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_float, C.c_int32, C.c_void_p]),
    "eags_image_l1_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "eags_image_l1_forward": (C.c_int, [C.POINTER(EagsImageL1), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_void_p]),
    "eags_image_l1_forward_static": (C.c_int, [C.POINTER(EagsImageL1), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p]),
    "eags_image_l1_backward": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p]),
    "eags_ssim_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "eags_ssim_forward": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "eags_ssim_backward": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p]),
    "eags_gaussian_activations": (C.c_int, [C.c_int32] + [C.c_void_p] * 7),
    "eags_mapping_loss_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "eags_mapping_loss_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "eags_mapping_loss_backward": (C.c_int, [C.c_void_p] * 8),
    "eags_isotropic_loss": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "eags_pose_compose": (C.c_int, [C.c_void_p] * 9),
    "eags_pose_chain": (C.c_int, [C.c_void_p] * 7),
    "eags_pose_se3_update": (C.c_int, [C.c_void_p] * 8),
    "eags_pose_adam_se3_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float,
                                            C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p]),
    "eags_neighbors_workspace_bytes": (C.c_size_t, [C.c_int64]),
    "eags_knn_mean_dist2": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "eags_radius_has_neighbor": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_float, C.c_void_p,
                                           C.c_void_p, C.c_size_t, C.c_void_p]),
    "eags_state_view": (C.c_int, [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int32,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p),
                                  C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
}

_lib = None


def load() -> C.CDLL:
    """Load the CUDA library; loud failure when it is missing (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the sm_100a CUDA library has not been built. Run "
            "`python -c 'import __graft_entry__ as g; g.build()'` at the repo root (or "
            "`make -C eags-slam_b200/csrc`). There is no CPU fallback for the rasterizer.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.eags_abi_version() != 2:
        raise ImportError("libeags_raster.so ABI version mismatch; rebuild it")
    _lib = lib
    return lib


def check(rc: int, where: str) -> None:
    if rc != 0:
        msg = load().eags_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{where} failed (code {rc}): {msg}")


def set_option(name: str, value: int) -> None:
    check(load().eags_set_option(name.encode(), int(value)), "eags_set_option")


def get_option(name: str) -> int:
    v = C.c_int32(0)
    check(load().eags_get_option(name.encode(), C.byref(v)), "eags_get_option")
    return int(v.value)


def kernel_times(max_spans: int = 256):
    """[(launcher name, device ms)] recorded by this thread since the last call (needs the
    "kernel_timing" option); synchronises with the recorded events."""
    names = (C.c_char_p * max_spans)()
    ms = (C.c_float * max_spans)()
    n = load().eags_kernel_times(max_spans, names, ms)
    return [(names[i].decode(), float(ms[i])) for i in range(n)]


def launch_count() -> int:
    return int(load().eags_launch_count())
