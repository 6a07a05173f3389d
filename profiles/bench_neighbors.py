"""Neighbour queries timed on the GPU with CUDA events, the CPU oracle beside them.
usage: python profiles/bench_neighbors.py [--reps 20]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

from eags_raster import scene  # noqa: E402
from eags_raster.neighbors import has_neighbor_within, mean_dist2_3nn  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn):
    for _ in range(3):
        fn()
    t = 0.0
    for _ in range(a.reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        t += e0.elapsed_time(e1)
    return t / a.reps


rows = []
for wl in ("c2", "c3", "c4"):
    p = scene.make_scene(wl, seed=0)["means3D"]
    n = p.shape[0]
    pg = p.cuda()
    ms = timed(lambda: mean_dist2_3nn(pg))
    # algorithmic bytes: read the cloud once to bin it (12 B), write + read the binned copy (16 + 16 B), 4 B out
    rows.append({"op": "distCUDA2", "workload": wl, "points": n, "ms": ms, "Mpoints_per_s": n / ms / 1e3,
                 "algorithmic_GBps": n * 48 / ms / 1e6, "hbm_frac": n * 48 / ms / 1e6 / peaks["hbm_gbs"]})
    if wl == "c2":
        from oracle import neighbors as nb
        t0 = time.perf_counter(); nb.mean_dist2_3nn(p.numpy()); rows[-1]["cpu_oracle_ms"] = (time.perf_counter() - t0) * 1e3
rng = np.random.default_rng(0)
ref = scene.make_scene("c2", seed=0)["means3D"].cuda()
qry = (ref[:50_000] + 0.05 * torch.randn(50_000, 3, device="cuda")).contiguous()
ms = timed(lambda: has_neighbor_within(ref, qry, 0.03))
rows.append({"op": "compute_new_points_ids", "ref_points": ref.shape[0], "query_points": qry.shape[0], "radius": 0.03, "ms": ms})
for r in rows:
    print(json.dumps(r))
