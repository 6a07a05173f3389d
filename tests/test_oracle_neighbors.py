"""CPU: the neighbour-query oracle (oracle/neighbors.py) -- its KD-tree shortcut
against its own brute force on clouds with duplicates / planar structure, and the faiss-style
radius rule of src/utils/mapper_utils.py:193-226."""
import numpy as np

from oracle import neighbors as nb


def clouds():
    rng = np.random.default_rng(3)
    uni = rng.uniform(-2, 2, (700, 3)).astype(np.float32)
    plane = uni.copy()
    plane[:, 2] = 0.25                                   # exactly planar
    dup = np.concatenate([uni[:300], uni[:50], uni[:50]])  # triplicated points: zero distances count
    return {"uniform": uni, "plane": plane, "duplicates": dup}


def test_kdtree_matches_bruteforce():
    for name, p in clouds().items():
        a, b = nb.mean_dist2_3nn_bruteforce(p), nb.mean_dist2_3nn(p)
        assert np.array_equal(a, b), name


def test_tiny_clouds_use_flt_max_like_upstream():
    p = np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0]], dtype=np.float32)
    fmax = np.finfo(np.float32).max
    out = nb.mean_dist2_3nn(p)
    assert np.allclose(out, (np.array([1 + 4, 1 + 5, 4 + 5], dtype=np.float32) + fmax) / 3, rtol=1e-6)
    assert np.isinf(nb.mean_dist2_3nn(p[:2])).all()     # FLT_MAX + FLT_MAX overflows, as upstream


def test_new_points_rule_is_squared_distance_below_radius():
    f = np.array([[0, 0, 0]], dtype=np.float32)
    q = np.array([[0.1, 0, 0], [0.2, 0, 0], [0.17, 0, 0], [0.18, 0, 0]], dtype=np.float32)
    ids, d = nb.new_points_ids(f, q, radius=0.03)       # sqrt(0.03) = 0.1732
    assert ids.tolist() == [1, 3]
    assert np.allclose(d, [0.01, 0.04, 0.0289, 0.0324], rtol=1e-5)
    ids, d = nb.new_points_ids(np.zeros((0, 3), np.float32), q)
    assert ids.tolist() == [0, 1, 2, 3] and d is None


def test_new_points_rule_against_reference_vectors():
    """tests/golden/neighbors_reference.npz: the reference's own compute_new_points_ids (mapper_utils.py:193-226)
    run over an exact brute-force stand-in for faiss.IndexFlatL2 (make_golden_neighbors.py)."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "neighbors_reference.npz"))
    for c in range(1, 6):
        ids, _ = nb.new_points_ids(g[f"{c}/ref"], g[f"{c}/new"], radius=float(g[f"{c}/radius"]))
        assert np.array_equal(ids, g[f"{c}/ids"]), c
