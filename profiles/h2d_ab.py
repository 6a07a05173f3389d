"""Interleaved A/B of the host->device copy of bench.py's e2e block (16.8 MB) from
  A: torch pin_memory()                      (cudaHostAlloc, 4 KiB pages)
  B: anonymous mmap + MADV_HUGEPAGE + cudaHostRegister   (transparent huge pages where the box has them)
with per-copy CUDA-event times, idle and while the SMs are busy, to see whether the 4x swings of the copy
rate on the shared boxes (profiles/r1c/e2e_h2d_variance.txt) depend on the host pages or on GPU load."""
import mmap
import os
import sys

import numpy as np
import torch

dev = torch.device("cuda:0")
torch.cuda.init()
N = int(16.8e6) // 4
HUGE = 2 << 20


def pinned_a():
    return torch.empty(N, dtype=torch.float32).pin_memory().fill_(1.0)


def pinned_b():
    size = ((N * 4 + HUGE - 1) // HUGE + 1) * HUGE
    m = mmap.mmap(-1, size, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
    try:
        m.madvise(mmap.MADV_HUGEPAGE)
    except Exception as e:                                  # noqa: BLE001
        print("madvise(MADV_HUGEPAGE) failed:", e)
    arr = np.frombuffer(m, dtype=np.float32)
    off = (-arr.ctypes.data % HUGE) // 4
    arr = arr[off:off + N]
    arr[:] = 1.0
    t = torch.from_numpy(arr)
    rc = torch.cuda.cudart().cudaHostRegister(t.data_ptr(), N * 4, 0)
    print("cudaHostRegister rc", rc, "is_pinned", t.is_pinned())
    return t, m


def thp_state():
    for f in ("/sys/kernel/mm/transparent_hugepage/enabled", "/sys/kernel/mm/transparent_hugepage/defrag"):
        try:
            print(f, open(f).read().strip())
        except OSError as e:
            print(f, e)
    try:
        for ln in open("/proc/self/smaps_rollup"):
            if "AnonHugePages" in ln:
                print(ln.strip())
    except OSError:
        pass


a = pinned_a()
b, _keep = pinned_b()
thp_state()
dst = [torch.empty(N, dtype=torch.float32, device=dev) for _ in range(2)]
cs = torch.cuda.Stream(device=dev)


def run(label, reps, busy):
    x = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16) if busy else None
    times = {"A": [], "B": []}
    evs = []
    if busy:
        for _ in range(reps // 2):                       # ~0.5 ms of tensor-core work per product
            x @ x
    with torch.cuda.stream(cs):
        for i in range(reps):
            for name, src, d in (("A", a, dst[0]), ("B", b, dst[1])):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(cs)
                d.copy_(src, non_blocking=True)
                e1.record(cs)
                evs.append((name, e0, e1))
    torch.cuda.synchronize()
    for name, e0, e1 in evs:
        times[name].append(e0.elapsed_time(e1))
    for name in ("A", "B"):
        t = np.array(times[name])
        q = np.percentile(t, [0, 10, 50, 90, 100])
        print(f"{label:5s} {name}: per-copy ms min {q[0]:.3f} p10 {q[1]:.3f} median {q[2]:.3f} p90 {q[3]:.3f} max {q[4]:.3f}"
              f"  -> median {N * 4 / q[2] / 1e6:.1f} GB/s", flush=True)
    return times


run("warm", 10, False)
for rnd in range(3):
    run("idle", 100, False)
    run("busy", 100, True)
