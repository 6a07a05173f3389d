"""GPU parity of the neighbour queries (csrc/knn.cu) against oracle/neighbors.py,
through the C ABI binding and the reference-level shims (`simple_knn._C.distCUDA2`,
`compute_new_points_ids` of src/utils/mapper_utils.py:193-226).

Tolerance: squared distances are fp32 sums of three products; kernel and oracle use the same
operation order, so the 3-NN means agree to 1e-6 relative (they are usually bit-identical)."""
import numpy as np
import pytest
import torch

from oracle import neighbors as nb

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def _clouds():
    rng = np.random.default_rng(11)
    uni = rng.uniform(-3, 3, (60_000, 3)).astype(np.float32)
    # surface-like: three noisy planes (what a depth camera sees) -- the cell size must adapt
    n = 40_000
    planes = []
    for ax in range(3):
        p = rng.uniform(0, 5, (n, 3)).astype(np.float32)
        p[:, ax] = 1.0 + 0.002 * rng.standard_normal(n).astype(np.float32)
        planes.append(p)
    surf = np.concatenate(planes)
    # two dense clusters far apart + isolated outliers (ring search cannot settle them)
    cl = np.concatenate([rng.normal(0, 0.05, (20_000, 3)), rng.normal(40, 0.05, (20_000, 3)),
                         rng.uniform(-500, 500, (40, 3))]).astype(np.float32)
    dup = np.concatenate([uni[:5000], uni[:700], uni[:700], uni[:700]])
    line = np.zeros((5000, 3), np.float32)
    line[:, 0] = np.linspace(0, 1, 5000, dtype=np.float32)   # degenerate extent on two axes
    return {"uniform": uni, "surfaces": surf, "clusters+outliers": cl, "duplicates": dup, "line": line}


@pytest.mark.parametrize("name", ["uniform", "surfaces", "clusters+outliers", "duplicates", "line"])
def test_dist_cuda2_matches_oracle(name):
    from simple_knn._C import distCUDA2
    p = _clouds()[name]
    got = distCUDA2(torch.from_numpy(p).cuda()).cpu().numpy()
    want = nb.mean_dist2_3nn(p)
    assert got.shape == want.shape and got.dtype == np.float32
    np.testing.assert_allclose(got, want, rtol=RTOL, atol=0)


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 33])
def test_dist_cuda2_tiny(n):
    from simple_knn._C import distCUDA2
    rng = np.random.default_rng(n)
    p = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    got = distCUDA2(torch.from_numpy(p).cuda()).cpu().numpy()
    want = nb.mean_dist2_3nn_bruteforce(p)
    assert np.array_equal(np.isinf(got), np.isinf(want))
    f = np.isfinite(want)
    np.testing.assert_allclose(got[f], want[f], rtol=RTOL)


def test_dist_cuda2_coincident_and_empty():
    from simple_knn._C import distCUDA2
    p = torch.full((1000, 3), 0.5, device="cuda")
    assert torch.count_nonzero(distCUDA2(p)) == 0
    assert distCUDA2(torch.zeros(0, 3, device="cuda")).shape == (0,)
    with pytest.raises(RuntimeError, match="CUDA"):
        distCUDA2(torch.zeros(4, 3))


def test_dist_cuda2_full_size_c2():
    """300 K points (the c2 cloud's means): every value against the oracle."""
    from eags_raster import scene
    from simple_knn._C import distCUDA2
    p = scene.make_scene("c2", seed=0)["means3D"].numpy()
    got = distCUDA2(torch.from_numpy(p).cuda()).cpu().numpy()
    want = nb.mean_dist2_3nn(p)
    np.testing.assert_allclose(got, want, rtol=RTOL, atol=0)


@pytest.mark.parametrize("radius", [0.03, 0.0004, 1.5])
def test_compute_new_points_ids_matches_oracle(radius):
    from eags_raster.neighbors import compute_new_points_ids
    rng = np.random.default_rng(5)
    frustum = rng.uniform(-2, 2, (150_000, 3)).astype(np.float32)
    new = np.concatenate([rng.uniform(-2.5, 2.5, (40_000, 3)),          # some outside the frustum box
                          frustum[:2000] + 1e-4,                        # near-duplicates
                          rng.uniform(50, 60, (100, 3))]).astype(np.float32)
    ids = compute_new_points_ids(torch.from_numpy(frustum).cuda(), torch.from_numpy(new).cuda(), radius=radius,
                                 device="cuda").cpu().numpy()
    want, d = nb.new_points_ids(frustum, new, radius=radius)
    # a query whose nearest squared distance sits within 1 ulp-ish of the radius may flip
    fragile = np.abs(d - np.float32(radius)) <= 1e-6 * radius
    got_mask = np.zeros(len(new), bool); got_mask[ids] = True
    want_mask = np.zeros(len(new), bool); want_mask[want] = True
    assert np.array_equal(got_mask[~fragile], want_mask[~fragile])
    assert ids.dtype == np.int64 and np.all(np.diff(ids) > 0)


def test_compute_new_points_ids_edge_cases():
    from eags_raster.neighbors import compute_new_points_ids
    new = torch.rand(10, 3, device="cuda")
    out = compute_new_points_ids(torch.zeros(0, 3, device="cuda"), new)
    assert out.tolist() == list(range(10))            # mapper_utils.py:207-208
    assert compute_new_points_ids(new, torch.zeros(0, 3, device="cuda")).numel() == 0
    # CPU inputs + device argument, as the reference's signature allows
    ids = compute_new_points_ids(new.cpu(), (new + 10.0).cpu(), radius=0.03, device="cuda")
    assert ids.tolist() == list(range(10))
    ids = compute_new_points_ids(new.cpu(), new.cpu(), radius=0.03, device="cuda")
    assert ids.numel() == 0


@pytest.mark.parametrize("c", range(1, 6))
def test_compute_new_points_ids_against_reference_vectors(c):
    """Kept indices equal the ones the reference's own compute_new_points_ids produced
    (tests/golden/neighbors_reference.npz, make_golden_neighbors.py; borderline queries were re-drawn there)."""
    import os
    from eags_raster.neighbors import compute_new_points_ids
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "neighbors_reference.npz"))
    ref, new = torch.from_numpy(g[f"{c}/ref"]).cuda(), torch.from_numpy(g[f"{c}/new"]).cuda()
    ids = compute_new_points_ids(ref, new, radius=float(g[f"{c}/radius"]), device="cuda")
    assert ids.dtype == torch.int64
    assert np.array_equal(ids.cpu().numpy(), g[f"{c}/ids"])
