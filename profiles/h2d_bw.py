"""Raw pinned-host -> device copy rate on this box for the e2e block of bench.py (c2: 16.8 MB), so the
e2e number can be read against the link it is bound by.  Prints GB/s for a pinned block allocated
(a) before and (b) after the process is pinned to the GPU-local CPUs, and for write-combined memory."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402

import bench  # noqa: E402

dev = torch.device("cuda:0")
torch.cuda.init()


def rate(host, dst, reps=30):
    s = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(s):
        for _ in range(3):
            dst.copy_(host, non_blocking=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            dst.copy_(host, non_blocking=True)
        e1.record(s)
    e1.synchronize()
    return host.numel() * 4 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9


for mb in (16.8, 67.2):
    n = int(mb * 1e6 / 4)
    dst = torch.empty(n, dtype=torch.float32, device=dev)
    h = torch.empty(n, dtype=torch.float32).pin_memory()
    print(f"{mb:5.1f} MB pinned (allocated before CPU pinning): {rate(h, dst):6.1f} GB/s", flush=True)
ncpu, before = bench.pin_to_gpu_numa_node(0)
print("GPU-local CPUs:", ncpu, "of", len(before) if before else None, flush=True)
for mb in (16.8, 67.2):
    n = int(mb * 1e6 / 4)
    dst = torch.empty(n, dtype=torch.float32, device=dev)
    h = torch.empty(n, dtype=torch.float32).pin_memory()
    print(f"{mb:5.1f} MB pinned (allocated after CPU pinning):  {rate(h, dst):6.1f} GB/s", flush=True)
    back = torch.empty(n, dtype=torch.float32).pin_memory()
    s = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(s):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        back.copy_(dst, non_blocking=True)
        e0.record(s)
        for _ in range(10):
            back.copy_(dst, non_blocking=True)
        e1.record(s)
    e1.synchronize()
    print(f"{mb:5.1f} MB D2H:                                    {n * 4 * 10 / (e0.elapsed_time(e1) * 1e-3) / 1e9:6.1f} GB/s", flush=True)
os.system("nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current,pcie.link.width.max --format=csv")
os.system("nvidia-smi topo -m | head -6")
