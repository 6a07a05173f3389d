"""Golden vectors for the radius test of keyframe seeding made by the reference's OWN
`compute_new_points_ids` (src/utils/mapper_utils.py:193-226), imported unchanged.  faiss is not installed here;
the function only needs `faiss.IndexFlatL2(3)` with add / search / reset, stood in for below by an exact
brute-force index with faiss' documented semantics (search returns the k smallest SQUARED L2 distances in
ascending order, missing neighbours padded with FLT_MAX / -1).  What the vectors pin is therefore the
reference's use of those results: 8 neighbours, `distances < radius` on squared distances, the empty-frustum
shortcut, the index dtype.  Queries whose nearest squared distance lies within 1e-3 relative of the radius are
re-drawn (faiss' BLAS formulation and a direct difference may round such a pair to different sides).

    python tests/golden/make_golden_neighbors.py   ->  tests/golden/neighbors_reference.npz
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


class IndexFlatL2:
    def __init__(self, d):
        self.d, self.pts = d, torch.zeros(0, d)

    def add(self, x):
        self.pts = torch.cat([self.pts, x.detach().float().cpu()], 0)

    def search(self, q, k):
        q = q.detach().float().cpu()
        d2 = ((q[:, None, :].double() - self.pts[None, :, :].double()) ** 2).sum(-1).float()
        n = self.pts.shape[0]
        if n < k:
            d2 = torch.cat([d2, torch.full((q.shape[0], k - n), torch.finfo(torch.float32).max)], 1)
        val, idx = torch.topk(d2, k, dim=1, largest=False, sorted=True)
        idx = torch.where(idx >= n, torch.full_like(idx, -1), idx)
        return val, idx

    def reset(self):
        self.pts = torch.zeros(0, self.d)


faiss = types.ModuleType("faiss")
faiss.IndexFlatL2 = IndexFlatL2
faiss.contrib = types.ModuleType("faiss.contrib")
faiss.contrib.torch_utils = types.ModuleType("faiss.contrib.torch_utils")
sys.modules.update({"faiss": faiss, "faiss.contrib": faiss.contrib, "faiss.contrib.torch_utils": faiss.contrib.torch_utils})
sys.path.insert(0, "/root/reference")
from src.utils.mapper_utils import compute_new_points_ids  # noqa: E402


def clouds(seed, n_ref, n_new, radius):
    g = torch.Generator().manual_seed(seed)
    ref = torch.rand(n_ref, 3, generator=g) * torch.tensor([2.0, 1.5, 1.0]) + torch.tensor([-1.0, -0.75, 1.0])
    # new points: half jittered copies of frustum points (mostly inside the radius), half fresh samples
    pick = torch.randint(0, max(n_ref, 1), (n_new // 2,), generator=g)
    jitter = torch.randn(n_new // 2, 3, generator=g) * (radius ** 0.5) * 0.8
    near = (ref[pick] + jitter) if n_ref else torch.rand(n_new // 2, 3, generator=g)
    fresh = torch.rand(n_new - n_new // 2, 3, generator=g) * torch.tensor([2.4, 1.9, 1.4]) + torch.tensor([-1.2, -0.95, 0.8])
    new = torch.cat([near, fresh], 0)
    if n_ref:
        for _ in range(20):                      # re-draw borderline queries
            d2 = ((new[:, None, :].double() - ref[None, :, :].double()) ** 2).sum(-1).min(1).values
            bad = (d2 / radius - 1).abs() < 1e-3
            if not bad.any():
                break
            new[bad] += torch.randn(int(bad.sum()), 3, generator=g) * 0.01
    return ref, new


def main():
    out = {}
    cases = [(1, 4000, 3000, 0.03), (2, 4000, 3000, 0.0009), (3, 5, 200, 0.03), (4, 0, 50, 0.03), (5, 20000, 1000, 0.001)]
    for seed, n_ref, n_new, radius in cases:
        ref, new = clouds(seed, n_ref, n_new, radius)
        ids = compute_new_points_ids(ref, new, radius=radius, device="cpu")
        assert ids.dtype == torch.int64
        out[f"{seed}/ref"], out[f"{seed}/new"] = ref.numpy(), new.numpy()
        out[f"{seed}/radius"] = np.float64(radius)
        out[f"{seed}/ids"] = ids.numpy()
        print(f"case {seed}: {n_ref} frustum points, {n_new} new, radius {radius}: {ids.numel()} kept")
    path = os.path.join(HERE, "neighbors_reference.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
