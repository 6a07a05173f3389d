"""Neighbour queries that bracket the rasterizer when a keyframe seeds new Gaussians
(SURVEY.md 8(f) N-3): drop-ins for the two un-vendored native dependencies the reference uses there.

* :func:`mean_dist2_3nn` == ``simple_knn._C.distCUDA2`` (reference call sites
  ``src/entities/gaussian_model.py:143,146,253,257,298,301,616``).
* :func:`compute_new_points_ids` mirrors ``src/utils/mapper_utils.py:193-226`` (faiss IndexFlatL2,
  8-NN search, ``distances < radius`` on SQUARED distances) -- same name, arguments, return value.

Both run on the hashed-grid kernels of ``csrc/knn.cu`` through the C ABI; there is no CPU fallback.
"""
from __future__ import annotations

import torch

from . import _lib


def _stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _points(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (the neighbour kernels have no CPU fallback)")
    if t.dim() != 2 or t.shape[1] != 3:
        raise ValueError(f"{name} must have shape (N, 3), got {tuple(t.shape)}")
    return t.detach().to(torch.float32).contiguous()


def mean_dist2_3nn(points: torch.Tensor) -> torch.Tensor:
    """(P,3) CUDA points -> (P,) mean squared distance to the 3 nearest other points."""
    pts = _points(points, "points")
    n = pts.shape[0]
    out = torch.empty(n, dtype=torch.float32, device=pts.device)
    if n == 0:
        return out
    lib = _lib.load()
    with torch.cuda.device(pts.device):
        nbytes = int(lib.eags_neighbors_workspace_bytes(n))
        ws = torch.empty(nbytes, dtype=torch.uint8, device=pts.device)
        _lib.check(lib.eags_knn_mean_dist2(pts.data_ptr(), n, out.data_ptr(), ws.data_ptr(), nbytes,
                                           _stream(pts.device)), "eags_knn_mean_dist2")
    return out


def has_neighbor_within(ref_points: torch.Tensor, query_points: torch.Tensor, radius: float) -> torch.Tensor:
    """(Q,) bool: query q has a reference point with squared distance < radius."""
    ref = _points(ref_points, "ref_points")
    qry = _points(query_points, "query_points")
    if ref.device != qry.device:
        raise ValueError("ref_points and query_points must live on the same device")
    nq, nr = qry.shape[0], ref.shape[0]
    mask = torch.zeros(nq, dtype=torch.uint8, device=qry.device)
    if nq == 0 or nr == 0:
        return mask.bool()
    lib = _lib.load()
    with torch.cuda.device(qry.device):
        nbytes = int(lib.eags_neighbors_workspace_bytes(nr))
        ws = torch.empty(nbytes, dtype=torch.uint8, device=qry.device)
        _lib.check(lib.eags_radius_has_neighbor(ref.data_ptr(), nr, qry.data_ptr(), nq, float(radius),
                                                mask.data_ptr(), ws.data_ptr(), nbytes, _stream(qry.device)),
                   "eags_radius_has_neighbor")
    return mask.bool()


def compute_new_points_ids(frustum_points: torch.Tensor, new_pts: torch.Tensor, radius: float = 0.03,
                           device: str = "cuda") -> torch.Tensor:
    """Indices of the new points with no frustum point within the radius
    (``mapper_utils.compute_new_points_ids``; the reference's ``device="cpu"`` default selects
    faiss-cpu -- here the search always runs on the GPU the points live on, or on ``device``)."""
    if frustum_points.shape[0] == 0:
        return torch.arange(new_pts.shape[0])
    dev = torch.device(device if device != "cpu" else "cuda") if not new_pts.is_cuda else new_pts.device
    mask = has_neighbor_within(frustum_points.to(dev), new_pts.to(dev), radius)
    return torch.where(~mask)[0]
