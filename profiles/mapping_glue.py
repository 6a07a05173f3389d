"""What runs between our kernels in one mapping step: the single-GPU c4 job (all 8 views) under torch.profiler,
every device activity (kernels, memsets, memcpys) summed per name and divided by the number of steps.

    python profiles/mapping_glue.py [--workload c4] [--steps 3] > gpurun_out/mapping_glue.txt
"""
import argparse
import os
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402

import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c4")
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--collective", default="peer")
ap.add_argument("--graph", action="store_true", help="profile CUDA-graph replays and list the idle gaps between device activities")
a = ap.parse_args()

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
job = bench.MappingJob(a.workload, dev, 1, 0, a.collective, group_world=False)
for _ in range(3):
    job.step()
torch.cuda.synchronize()

# (a) eager wall / device time per step
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for _ in range(a.steps):
    job.step()
ev[1].record()
torch.cuda.synchronize()
print(f"eager: {ev[0].elapsed_time(ev[1]) / a.steps:.3f} ms per step ({bench.N_VIEWS} views)")

from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(a.steps):
        job.step()
    torch.cuda.synchronize()

tot, cnt = defaultdict(float), defaultdict(int)
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        tot[e.name] += e.device_time if hasattr(e, "device_time") else e.cuda_time
        cnt[e.name] += 1
ours = ("eags", "preprocess", "blend", "sort_pack", "tile_scan", "duplicate", "image_l1", "ssim", "peer_", "horizontal",
        "vertical")
rows = sorted(tot.items(), key=lambda kv: -kv[1])
s_all = sum(tot.values())
s_ours = sum(v for k, v in tot.items() if any(o in k for o in ours))
print(f"device activity per step: {s_all / a.steps / 1e3:.3f} ms, of it our kernels {s_ours / a.steps / 1e3:.3f} ms, "
      f"other {(s_all - s_ours) / a.steps / 1e3:.3f} ms")
print(f"{'us/step':>10} {'n/step':>7}  name")
for k, v in rows:
    tag = " " if any(o in k for o in ours) else "*"
    print(f"{v / a.steps:10.1f} {cnt[k] / a.steps:7.1f} {tag} {k[:150]}")

if a.graph:
    # the same step replayed from a CUDA graph: where is the device idle between two activities?
    from eags_raster.graph import GraphedIteration
    gi = GraphedIteration(job.step, warmup=1, max_renders=2 * bench.N_VIEWS)
    for _ in range(3):
        gi.replay()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(a.steps):
        gi.replay()
    ev[1].record()
    torch.cuda.synchronize()
    print(f"\ngraph replay: {ev[0].elapsed_time(ev[1]) / a.steps:.3f} ms per step")
    with profile(activities=[ProfilerActivity.CUDA]) as prof2:
        for _ in range(a.steps):
            gi.replay()
        torch.cuda.synchronize()
    iv = sorted((e.time_range.start, e.time_range.end, e.name) for e in prof2.events()
                if e.device_type == torch.autograd.DeviceType.CUDA)
    busy = sum(e - s_ for s_, e, _ in iv)
    span = iv[-1][1] - iv[0][0]
    gaps = defaultdict(lambda: [0.0, 0])
    end, last = iv[0][1], iv[0][2]
    for s_, e, n in iv[1:]:
        if s_ > end:
            g = gaps[(last[:48], n[:48])]
            g[0] += s_ - end
            g[1] += 1
        if e > end:
            end, last = e, n
    idle = sum(g[0] for g in gaps.values())
    print(f"under the profiler, {a.steps} replays: span {span / a.steps / 1e3:.3f} ms/step, busy {busy / a.steps / 1e3:.3f}, idle {idle / a.steps / 1e3:.3f}")
    print(f"{'us/step':>9} {'n/step':>7}  gap after -> before")
    for (p_, n_), (t, c) in sorted(gaps.items(), key=lambda kv: -kv[1][0])[:25]:
        print(f"{t / a.steps:9.1f} {c / a.steps:7.1f}  {p_} -> {n_}")
