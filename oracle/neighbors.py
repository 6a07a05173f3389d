"""ORACLE (test infrastructure, never imported by the product): CPU restatement of the two neighbour
queries of SURVEY.md 8(f) N-3.

* ``mean_dist2_3nn`` follows upstream simple-knn (``simple_knn.cu``: ``updateKBest<3>`` /
  ``boxMeanDist``; pinned by the reference only through its call sites,
  ``src/entities/gaussian_model.py:143-146``): for every point the three smallest squared distances
  to OTHER points (self excluded by index, duplicates count), fp32, ``(b0 + b1 + b2) / 3``.
  PARITY UNPINNED: simple-knn is an un-vendored pip dependency; the reference holds no test or
  golden vector for it.
* ``new_points_ids`` follows ``src/utils/mapper_utils.py:193-226``: keep the new points whose nearest
  frustum point has squared distance >= radius (faiss IndexFlatL2 returns squared distances).
"""
import numpy as np


def _d2_f32(a, b):
    """fp32 squared distance in the kernel's operation order (dx*dx, then two fused multiply-adds)."""
    d = (a.astype(np.float32) - b.astype(np.float32)).astype(np.float32)
    dx, dy, dz = (d[..., k].astype(np.float64) for k in range(3))
    t = (dx * dx).astype(np.float32).astype(np.float64)      # rounded product
    t = (dy * dy + t).astype(np.float32).astype(np.float64)  # fma: exact product + add, one rounding
    return (dz * dz + t).astype(np.float32)


def mean_dist2_3nn_bruteforce(points):
    p = np.asarray(points, dtype=np.float32)
    n = p.shape[0]
    out = np.empty(n, dtype=np.float32)
    fmax = np.float32(np.finfo(np.float32).max)
    for i in range(n):
        d = _d2_f32(p, p[i][None, :])
        d[i] = np.inf
        k = np.sort(d)[:3]
        best = np.full(3, fmax, dtype=np.float32)
        best[:len(k)] = np.where(np.isinf(k), fmax, k)
        with np.errstate(over="ignore"):
            out[i] = (best[0] + best[1] + best[2]) / np.float32(3.0)
    return out


def mean_dist2_3nn(points):
    """KD-tree candidates (float64) + fp32 distances in the kernel's order; needs >= 4 points."""
    from scipy.spatial import cKDTree
    p = np.asarray(points, dtype=np.float32)
    n = p.shape[0]
    if n < 8:
        return mean_dist2_3nn_bruteforce(p)
    k = min(n, 8)   # a few extra candidates: fp32 rounding may reorder near-ties
    _, idx = cKDTree(p.astype(np.float64)).query(p.astype(np.float64), k=k)
    out = np.empty(n, dtype=np.float32)
    d = np.stack([_d2_f32(p[idx[:, j]], p) for j in range(k)], axis=1)
    self_col = idx == np.arange(n)[:, None]
    # exactly one self entry must be dropped per row (duplicates of the point stay)
    first_self = np.argmax(self_col, axis=1)
    has_self = self_col.any(axis=1)
    d[np.arange(n)[has_self], first_self[has_self]] = np.inf
    d.sort(axis=1)
    out[:] = (d[:, 0] + d[:, 1] + d[:, 2]) / np.float32(3.0)
    return out


def new_points_ids(frustum_points, new_pts, radius=0.03):
    """-> (kept indices, fp32 squared distance to the nearest frustum point or None)."""
    from scipy.spatial import cKDTree
    f = np.asarray(frustum_points, dtype=np.float32)
    q = np.asarray(new_pts, dtype=np.float32)
    if f.shape[0] == 0:
        return np.arange(q.shape[0]), None
    _, idx = cKDTree(f.astype(np.float64)).query(q.astype(np.float64), k=min(4, f.shape[0]))
    idx = idx.reshape(q.shape[0], -1)
    d = np.stack([_d2_f32(f[idx[:, j]], q) for j in range(idx.shape[1])], axis=1).min(axis=1)
    return np.nonzero(~(d < np.float32(radius)))[0], d
