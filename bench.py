#!/usr/bin/env python
"""bench.py -- rasterizer fwd+bwd iterations/s on the BASELINE.json workloads.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2] [--impl ours|reference]

One *step* = one pass of the hot path over one synthetic view: GaussianRasterizer forward ->
L1 colour + L1 depth loss -> backward to means3D / opacities / scales / rotations.
N = 1 runs BASELINE.json configs[1] (c2: 640x480, 300 K Gaussians, the tracking shape).  N > 1 is
the sharded-mapping shape: every rank holds a replica of the same submap, renders its own keyframe
view per step, and the per-Gaussian gradients (11 floats) are summed with one NCCL all-reduce
(weak scaling: the per-GPU work is the N = 1 workload; value = view-iterations/s over all ranks).

Timing: W >= 3 warm-up steps, then K steps each bracketed by CUDA events on the launching stream
with an L2 flush (256 MiB write) before every step, outside the events; barrier +
torch.cuda.synchronize() on both sides; max over ranks.  `e2e` repeats the measurement through
the same public API with HOST (pinned) inputs: host->device copies of all Gaussian tensors and a
device->host read of the loss inside every timed step.  `--impl reference` times the CPU oracle
(the reference has no CPU implementation of this path and its CUDA kernels are un-vendored pip
dependencies) on the box's host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "eags-slam_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

METRIC = "rasterizer fwd+bwd it/s"
UNIT = "it/s"
FP32_FLOPS_FWD_PAIR = 30.0
FP32_FLOPS_BWD_PAIR = 90.0


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm_gbs=float(d["hbm_gbs"]), sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)), src="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        if os.environ.get("EAGS_BENCH_CLOCKS", "smi") == "off":      # A/B knob: does the polling perturb the run?
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index),
                 "-lms", os.environ.get("EAGS_BENCH_CLOCKS_MS", "20")], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def make_workload(name, device, seed, view_id=0, P=None):
    from eags_raster import scene
    from gaussian_rasterizer import GaussianRasterizationSettings
    sc = scene.make_scene(name, P=P, seed=seed)
    if view_id == 0:
        cam = scene.workload_camera(name)
    else:
        cam = scene.keyframe_views(name, 8)[view_id % 8]
    s = GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=torch.zeros(3, device=device), scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(device),
        projmatrix=cam.projmatrix.to(device), sh_degree=0, campos=cam.campos.to(device),
        prefiltered=False, debug=False)
    return sc, cam, s


_T0 = time.perf_counter()


def _stamp(what):
    if os.environ.get("EAGS_BENCH_STAMPS"):
        print(f"[bench +{time.perf_counter() - _T0:7.2f}s] {what}", file=sys.stderr, flush=True)


def pin_to_gpu_numa_node(index):
    """Run this process (and allocate its pinned host buffers) on the CPUs NVML reports as local to the
    GPU: host<->device copies from the far socket run at about half the PCIe rate."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1}
        before = os.sched_getaffinity(0)
        cpus &= before
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus), before
    except Exception:
        pass
    return None, None


_HOST_BLOCKS = []     # (mmap, pointer) of huge-page staging blocks: kept alive for the life of the process


def pinned_host_block(numel):
    """Pinned fp32 staging block for the e2e leg.  Anonymous mmap + MADV_HUGEPAGE + cudaHostRegister: alone on a
    B200 box the first ~10 copies out of a 4-KiB-page cudaHostAlloc block (torch ``pin_memory()``) take 0.5 - 8 ms
    per 16.8 MB while the huge-page block streams at 54 GB/s from the first copy on (profiles/r1c/h2d_ab.txt);
    in steady state both are equal, and inside bench.py the A/B shows no difference (the slow phases of
    profiles/r1c/e2e_h2d_variance.txt hit both).  EAGS_BENCH_HOSTMEM=torch, or any failure, falls back to
    ``pin_memory()``.  Returns (tensor, kind)."""
    if os.environ.get("EAGS_BENCH_HOSTMEM", "hugepage") != "torch":
        try:
            import mmap
            import numpy as np
            huge = 2 << 20
            size = ((numel * 4 + huge - 1) // huge + 1) * huge
            m = mmap.mmap(-1, size, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
            m.madvise(mmap.MADV_HUGEPAGE)
            arr = np.frombuffer(m, dtype=np.float32)
            off = (-arr.ctypes.data % huge) // 4
            arr = arr[off:off + numel]
            arr[:] = 0.0                                            # touch: populate the (huge) pages
            t = torch.from_numpy(arr)
            rc = torch.cuda.cudart().cudaHostRegister(t.data_ptr(), numel * 4, 0)
            if int(rc) == 0 and t.is_pinned():
                _HOST_BLOCKS.append((m, t.data_ptr()))
                return t, "mmap + MADV_HUGEPAGE + cudaHostRegister"
        except Exception:                                           # noqa: BLE001
            pass
    return torch.empty(numel, dtype=torch.float32).pin_memory(), "torch pin_memory (cudaHostAlloc)"


def run_ours(args):
    import torch.distributed as dist
    from eags_raster import _lib, rasterize as rz, state_array
    from eags_raster.losses import masked_l1_means
    from gaussian_rasterizer import GaussianRasterizer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_cpus, all_cpus = pin_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()
    peaks = load_peaks()
    wl = args.workload
    # every rank: the SAME submap (seed 0), its own keyframe view
    sc, cam, settings = make_workload(wl, dev, seed=0, view_id=rank if world > 1 else 0)
    keys = ("means3D", "scales", "rotations", "opacities")
    P = sc["means3D"].shape[0]
    host = {k: sc[k].pin_memory() for k in keys + ("shs",)}
    devt = {k: sc[k].to(dev) for k in keys + ("shs",)}
    gt_c, gt_d = sc["gt_color"].to(dev), sc["gt_depth"].to(dev)
    raster = GaussianRasterizer(raster_settings=settings)
    # sharded mapping (N > 1): p.grad of every leaf is a view into ONE fused buffer of 11 floats /
    # Gaussian (xyz 3, scale 3, rot 4, opacity 1) -> one NCCL all-reduce per step, nothing packed
    sizes = {"means3D": 3, "scales": 3, "rotations": 4, "opacities": 1}
    flat_grad = torch.zeros(P * 11, device=dev)
    grad_views, off = {}, 0
    for k in keys:
        grad_views[k] = flat_grad[off:off + P * sizes[k]].view(P, sizes[k])
        off += P * sizes[k]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    # N > 1 gradient exchange: "peer" = reduce-scatter fused into preprocess_bwd + one peer_reduce kernel
    # over NVLink peer memory (eags_raster.peer, csrc/collective.cu); "nccl" = backward into the fused
    # bucket + one NCCL all-reduce (the baseline; also the fallback when peer memory is unavailable)
    exchange, collective = None, "none"
    if world > 1:
        collective = args.collective
        if collective == "peer":
            try:
                from eags_raster.peer import PeerGradExchange
                exchange = PeerGradExchange(P, dev)
            except Exception as e:   # no peer access / symmetric memory on this box: say so, use NCCL
                collective = f"nccl (peer exchange unavailable: {type(e).__name__}: {str(e)[:120]})"
                flag = torch.ones(1, device=dev)
            else:
                flag = torch.zeros(1, device=dev)
            dist.all_reduce(flag)    # all ranks must take the same path
            if float(flag) > 0 and exchange is not None:
                exchange, collective = None, "nccl (peer exchange unavailable on another rank)"

    def step(src, read_loss=False):
        leaf = {k: src[k].detach().requires_grad_(True) for k in keys}
        m2d = torch.zeros_like(leaf["means3D"], requires_grad=True)
        color, depth, alpha, radii = raster(means3D=leaf["means3D"], means2D=m2d, opacities=leaf["opacities"],
                                            shs=src["shs"], colors_precomp=None, scales=leaf["scales"],
                                            rotations=leaf["rotations"], cov3D_precomp=None)
        if args.loss == "fused":    # the product's fused loss kernels (eags_raster.losses)
            lc, ld = masked_l1_means(color, depth, gt_c, gt_d)
            loss = lc + ld
        else:                       # the same loss from stock torch ops
            loss = (color - gt_c).abs().mean() + (depth - gt_d).abs().mean()
        if exchange is not None:
            # gradients leave preprocess_bwd through NVLink to their owner rank; p.grad comes back
            # summed over all ranks (x 1/world), bit-identical on every replica
            with exchange.route(scale=1.0 / world):
                loss.backward()
        elif world > 1:
            # the backward kernel writes the 11 floats / Gaussian straight into the fused buffer
            with rz.grad_arena(**grad_views):
                loss.backward()
            dist.all_reduce(flat_grad)
        else:
            loss.backward()
        return float(loss.detach()) if read_loss else loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    _stamp("setup done")
    # ---- warm-up, then device-timed steps with inputs resident in HBM ----
    for _ in range(max(args.warmup, 3)):
        step(devt)
    barrier()
    # launch mode "graph" (N = 1): the iteration built from the public API is captured ONCE into a CUDA
    # graph (eags_raster.graph.GraphedIteration; static input / gradient tensors, as a tracking or mapping
    # loop has) and replayed -- the same kernels, one launch per step instead of ~0.4 ms of Python, autograd
    # and launch work plus one host wait.  "eager" issues every step through Python.
    use_graph = args.launch == "graph"
    order = keys + ("shs",)
    sizes_h = [devt[k].numel() for k in order]

    def views_of(flat):
        bufs, o = {}, 0
        for k, n_ in zip(order, sizes_h):
            bufs[k] = flat[o:o + n_].view(devt[k].shape)
            o += n_
        return bufs

    gi, graph_launches, graph_error = None, None, None
    if use_graph:
        from eags_raster.graph import GraphedIteration
        static_flat = torch.cat([devt[k].reshape(-1) for k in order])
        static = views_of(static_flat)
        leaves = {k: static[k].requires_grad_(True) for k in keys}

        def graph_fn():
            for t_ in leaves.values():
                t_.grad = None
            m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
            color, depth, alpha, radii = raster(means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"],
                                                shs=static["shs"], colors_precomp=None, scales=leaves["scales"],
                                                rotations=leaves["rotations"], cov3D_precomp=None)
            if args.loss == "fused":
                lc, ld = masked_l1_means(color, depth, gt_c, gt_d)
                loss = lc + ld
            else:
                loss = (color - gt_c).abs().mean() + (depth - gt_d).abs().mean()
            if exchange is not None:
                with exchange.route(scale=1.0 / world):
                    loss.backward()
            elif world > 1:
                with rz.grad_arena(**grad_views):
                    loss.backward()
                dist.all_reduce(flat_grad)
            else:
                loss.backward()
            return loss.detach()

        lc0 = _lib.launch_count()
        try:
            gi = GraphedIteration(graph_fn, warmup=1)
            ok = torch.ones(1, device=dev)
        except Exception as e:       # e.g. a collective that cannot be captured on this stack
            gi, ok = None, torch.zeros(1, device=dev)
            graph_error = f"{type(e).__name__}: {str(e)[:160]}"
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)    # all ranks replay, or none does
        if float(ok) < 1:
            use_graph, gi = False, None
        else:
            graph_launches = (_lib.launch_count() - lc0) / 2.0     # one eager warm-up + the captured pass
            for _ in range(3):
                gi.replay()
            gi.check()
    _stamp("graph captured" if use_graph else "eager mode")
    device_step = (lambda: gi.replay()) if use_graph else (lambda: step(devt))
    l0 = _lib.launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    t_wall0 = time.perf_counter()
    for a, b in evs:
        flush.zero_()
        a.record()
        device_step()
        b.record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    _stamp("timed steps done")
    launches = graph_launches if use_graph else (_lib.launch_count() - l0) / args.steps
    ms = sum(a.elapsed_time(b) for a, b in evs) / args.steps
    eager = None
    if use_graph:
        gi.check()     # no replay outgrew the captured buffers (it would not have been rendered)
        n_e = min(args.steps, 100)
        evs_e = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_e)]
        for a, b in evs_e:
            flush.zero_()
            a.record()
            step(devt)
            b.record()
        barrier()
        ms_e = sum(a.elapsed_time(b) for a, b in evs_e) / n_e
        if world > 1:
            t_e = torch.tensor([ms_e], device=dev, dtype=torch.float64)
            dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
            ms_e = float(t_e[0])
        eager = {"value": world * 1e3 / ms_e, "unit": UNIT, "ms_per_step": ms_e, "steps": n_e,
                 "note": "the same step issued eagerly through Python every iteration (sync_mode optimistic)"}
    # ---- e2e: host (pinned) inputs, H2D of every Gaussian tensor + D2H of the loss, per step ----
    # Input pipeline: the copy of step i+1's inputs is issued on a copy stream while step i computes
    # (double buffering, as a data loader would); every step's inputs cross PCIe inside the timed
    # region and the loss is read back (host sync) every step.
    # All five Gaussian tensors live in ONE pinned host block and cross PCIe as one copy per step (56 B /
    # Gaussian: 0.31 ms at the link's 54 GB/s, shorter than the kernels; up to 1.4 ms when the host side of
    # the shared box is busy, and then the step is copy-bound -- see link_rate below); the copies run two
    # steps ahead on a copy stream so the link never idles while the host reads the loss back.
    host_flat, hostmem_kind = pinned_host_block(sum(sizes_h))
    o = 0
    for k, n_ in zip(order, sizes_h):
        host_flat[o:o + n_].copy_(host[k].reshape(-1))
        o += n_
    h2d_bytes = host_flat.numel() * 4
    copy_stream = torch.cuda.Stream(device=dev)

    # static ring of device buffers (no allocator traffic inside the timed region)
    RING = 3
    ring = [torch.empty(host_flat.numel(), dtype=torch.float32, device=dev) for _ in range(RING)]
    ring_free = [None] * RING   # event: the step that last read ring[j] has been issued and finished

    ring_views = [views_of(f) for f in ring]
    loss_host = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]

    # EAGS_BENCH_E2E_SPLIT=k splits the block over k copy streams (A/B knob; one stream is the default)
    nsplit = max(1, int(os.environ.get("EAGS_BENCH_E2E_SPLIT", "1")))
    copy_streams = [copy_stream] + [torch.cuda.Stream(device=dev) for _ in range(nsplit - 1)]
    bounds = [host_flat.numel() * c // nsplit for c in range(nsplit + 1)]
    copy_times = [] if os.environ.get("EAGS_BENCH_E2E_TRACE") else None

    def prefetch(j):
        evs = []
        for c, cs in enumerate(copy_streams):
            with torch.cuda.stream(cs):
                if ring_free[j] is not None:
                    cs.wait_event(ring_free[j])
                if copy_times is not None and c == 0:
                    t_a = torch.cuda.Event(enable_timing=True)
                    t_a.record(cs)
                ring[j][bounds[c]:bounds[c + 1]].copy_(host_flat[bounds[c]:bounds[c + 1]], non_blocking=True)
                ev = torch.cuda.Event(enable_timing=copy_times is not None and c == 0)
                ev.record(cs)
                if copy_times is not None and c == 0:
                    copy_times.append((t_a, ev))
            evs.append(ev)
        return evs

    def e2e_loop(n, depth=2):
        for j in range(RING):
            ring_free[j] = None
        q = [prefetch(i % RING) for i in range(min(depth, n))]
        pending, losses_read = None, []
        trace = [] if os.environ.get("EAGS_BENCH_E2E_TRACE") else None
        for i in range(n):
            ev = q.pop(0)
            if i + depth < n:
                q.append(prefetch((i + depth) % RING))
            cur = torch.cuda.current_stream()
            for e_ in ev:
                cur.wait_event(e_)
            if use_graph:     # this step's inputs: ring slot -> the graph's static input block (D2D)
                static_flat.copy_(ring[i % RING], non_blocking=True)
                loss = gi.replay()
            else:
                loss = step(ring_views[i % RING])
            # the loss crosses to the host every step (4 B D2H into pinned memory); the host consumes it
            # one step later, as a training loop that logs its loss does, so the read never drains the GPU
            loss_host[i % 2].copy_(loss.detach(), non_blocking=True)
            done = torch.cuda.Event()
            done.record(cur)
            ring_free[i % RING] = done
            if pending is not None:
                pending[0].synchronize()
                losses_read.append(float(pending[1]))
                if trace is not None:
                    trace.append(time.perf_counter())
            pending = (done, loss_host[i % 2])
        if pending is not None:
            pending[0].synchronize()
            losses_read.append(float(pending[1]))
        assert len(losses_read) == n and all(v == v for v in losses_read)
        if trace:
            print("[e2e trace] ms between consecutive step completions:",
                  " ".join(f"{(b - a) * 1e3:.2f}" for a, b in zip(trace[:40], trace[1:41])), file=sys.stderr, flush=True)
            torch.cuda.synchronize()
            d = sorted(a.elapsed_time(b) for a, b in copy_times)
            print(f"[e2e trace] H2D copy on stream 0 ({(bounds[1] - bounds[0]) * 4 / 1e6:.1f} MB of {nsplit} split(s)): "
                  f"min {d[0]:.3f} median {d[len(d) // 2]:.3f} max {d[-1]:.3f} ms", file=sys.stderr, flush=True)
            copy_times.clear()

    def link_rate(reps=10):
        """The same block, copied alone (nothing else queued), CUDA events on the copy stream: the host->device
        rate the box delivers right now.  The e2e step is max(copy, compute); on the shared boxes the copy moves
        between 0.31 and 1.4 ms (profiles/r1c/e2e_h2d_variance.txt), so every e2e number carries this figure."""
        try:
            torch.cuda.synchronize()
            with torch.cuda.stream(copy_stream):
                pairs = []
                for _ in range(reps):
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(copy_stream)
                    ring[0].copy_(host_flat, non_blocking=True)
                    b.record(copy_stream)
                    pairs.append((a, b))
            copy_stream.synchronize()
            t = sorted(a.elapsed_time(b) for a, b in pairs)
            return {"median_ms": t[len(t) // 2], "min_ms": t[0], "max_ms": t[-1],
                    "median_gbs": h2d_bytes / t[len(t) // 2] / 1e6}
        except Exception as e:                                      # noqa: BLE001
            return {"error": f"{type(e).__name__}: {str(e)[:100]}"}

    _stamp("eager comparison done")
    e2e_loop(20)      # the host pipeline needs ~20 steps to settle (per-step trace: EAGS_BENCH_E2E_TRACE=1)
    link_before = link_rate()
    barrier()
    t0 = time.perf_counter()
    e2e_loop(args.steps)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    link_after = link_rate()
    if use_graph:
        gi.check()
    _stamp("e2e done")
    clocks = sampler.stop() if rank == 0 else None
    # ---- max over ranks ----
    if world > 1:
        t = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms = float(t[0]), float(t[1])

    out = None
    if rank == 0:
        # per-kernel timing of one more forward/backward with events around each C-ABI stage
        roof, kern = kernel_rooflines(wl, settings, devt, gt_c, gt_d, peaks, rz, state_array)
        out = {
            "metric": METRIC, "value": world * 1e3 / ms, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{wl}: {cam.W}x{cam.H}, {P} Gaussians, fwd + L1 colour/depth loss + bwd",
                       "views_per_step_per_gpu": 1, "loss": args.loss,
                       "launch": ("cuda-graph replay of the iteration captured from the public API (eags_raster.graph)"
                                  if use_graph else "eager (every step issued through Python)" + (f"; graph capture failed: {graph_error}" if graph_error else "")), "sync_mode": rz.get_sync_mode(), "tma": _lib.get_option("tma"),
                       "l2": "flushed before every timed step (256 MiB write, outside the events)",
                       "parallelism": "1 view/GPU, replica per GPU" + (", gradient exchange of 11 floats/Gaussian: " + collective if world > 1 else "")},
            "e2e": {"value": world * 1e3 / e2e_ms, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms,
                    "host_block": hostmem_kind,
                    "h2d_copy_alone": {"before": link_before, "after": link_after,
                                       "note": "the step's block copied alone right before / after the e2e leg"},
                    "pipeline": "one pinned block -> one H2D copy per step, issued two steps ahead on a copy stream into a static ring of device buffers; every step's loss is copied D2H and consumed by the host one step later"},
            # kernels of this repo launched inside the timed region (counted by the C ABI's launch counter:
            # per step x steps; under graph replay each replay launches the captured kernels once)
            "gpu_launches": int(round(launches * args.steps)), "gpu_launches_per_step": launches, "eager": eager, "host_cpus_local_to_gpu": numa_cpus,
            "wall_ms_per_step_incl_flush": t_wall * 1e3 / args.steps,
            "clocks": clocks, "roofline": roof, "kernels": kern,
        }
        if all_cpus:
            os.sched_setaffinity(0, all_cpus)     # the CPU baseline uses every host core
        if world == 1 and not args.no_cpu:
            out["cpu_baseline"] = cpu_baseline(wl, budget_s=args.cpu_budget)
        print(json.dumps(out), flush=True)
    _stamp("json printed")
    if world > 1:
        dist.barrier()
        _stamp("final barrier passed")
        dist.destroy_process_group()
        _stamp("process group destroyed")
    return out


# bound and algorithmic work of every launcher
def algorithmic_work(P, K, N, S_fwd, S_bwd):
    return {
        "preprocess_fwd": ("hbm", P * 128),
        "tile_scan": ("latency", None),
        "duplicate": ("hbm", P * 20 + K * 8),
        "sort_pack": ("hbm", K * (8 + 40 + 52)),
        "blend_fwd": ("fp32", S_fwd * FP32_FLOPS_FWD_PAIR),
        "blend_bwd": ("fp32", S_bwd * FP32_FLOPS_BWD_PAIR),
        "preprocess_bwd": ("hbm", P * 250),
    }


def ncu_traffic(wl, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed ncu capture."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(wl, {}).get(kernel)
    except Exception:
        return None


def kernel_rooflines(wl, settings, devt, gt_c, gt_d, peaks, rz, state_array, reps=20):
    """Live per-kernel durations (CUDA events recorded by the library on the launching stream
    around each launcher -- the "kernel_timing" option of the C ABI -- with the L2 flushed before
    every forward and every backward) and the roofline of each kernel; `roofline` is the dominant
    one."""
    from eags_raster import _lib
    dev = devt["means3D"].device
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    H, W = settings.image_height, settings.image_width
    dC = torch.sign(torch.randn(3, H, W, device=dev)) / (3 * H * W)
    dD = torch.sign(torch.randn(1, H, W, device=dev)) / (H * W)
    tf = tb = 0.0
    per = {}
    _lib.set_option("kernel_timing", 1)
    try:
        for i in range(reps + 3):
            flush.zero_()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            e[0].record()
            out = rz.rasterize_forward(settings, 0, devt["means3D"], devt["opacities"], shs=devt["shs"],
                                       scales=devt["scales"], rotations=devt["rotations"])
            e[1].record()
            st = out[5]
            flush.zero_()
            e2 = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            e2[0].record()
            rz.rasterize_backward(settings, st, devt["means3D"], devt["opacities"], devt["shs"], None,
                                  devt["scales"], devt["rotations"], None, dC, dD, None, want_shs=False)
            e2[1].record()
            torch.cuda.synchronize()
            spans = _lib.kernel_times()
            if i >= 3:
                tf += e[0].elapsed_time(e[1])
                tb += e2[0].elapsed_time(e2[1])
                for name, ms in spans:
                    per[name] = per.get(name, 0.0) + ms
    finally:
        _lib.set_option("kernel_timing", 0)
    tf, tb = tf / reps, tb / reps
    per = {k: v / reps for k, v in per.items()}
    K = st.num_rendered
    P = st.P
    N = H * W
    ncontrib = state_array(st, "n_contrib").to(torch.int64)
    ranges = state_array(st, "ranges").to(torch.int64)
    gx = (W + 15) // 16
    tile_n = (ranges[:, 1] - ranges[:, 0])
    ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
    per_pix_n = tile_n[: ((H + 15) // 16) * gx][(ys // 16) * gx + xs // 16]
    final_T = state_array(st, "final_T")
    # S: list entries a pixel iterates before it stops; saturated pixels stop early
    # (bounded below by their n_contrib), the others walk their whole tile list.
    S = int(torch.where(final_T * 0.01 < 1e-4, ncontrib, per_pix_n).sum())
    S_bwd = int(ncontrib.sum())
    fp32_peak = 148 * 128 * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12
    work = algorithmic_work(P, K, N, S, S_bwd)
    table = {}
    for name, ms in per.items():
        bound, amount = work.get(name, ("latency", None))
        row = {"ms": ms, "bound": bound}
        if bound == "hbm":
            row.update(achieved=amount / (ms * 1e-3) / 1e9, peak=peaks["hbm_gbs"], unit="GB/s",
                       algorithmic_bytes=amount)
        elif bound == "fp32":
            row.update(achieved=amount / (ms * 1e-3) / 1e12, peak=fp32_peak, unit="TFLOP/s",
                       algorithmic_flops=amount)
        if "achieved" in row:
            row["frac"] = row["achieved"] / row["peak"]
        row["traffic"] = ncu_traffic(wl, name)
        row["ncu"] = ncu_traffic(wl + "_ncu", name)   # issue-slot / pipe utilisation of the committed ncu capture
        table[name] = row
    top = max(table, key=lambda k: table[k]["ms"])
    t = table[top]
    roof = {
        "kernel": top, "bound": t["bound"], "achieved": t.get("achieved"), "peak": t.get("peak"),
        "unit": t.get("unit"), "frac": t.get("frac"), "traffic": t["traffic"], "ms": t["ms"], "ncu": t.get("ncu"),
        "peak_source": (f"fp32: 148 SM x 128 lanes x 2 x {peaks['sm_max_mhz']:.0f} MHz; hbm: {peaks['hbm_gbs']:.0f} GB/s "
                        f"(MEASURED_PEAKS.json, {peaks['src']})"),
        "note": "blending is FP32-pipe / issue-slot bound (no tensor-core work on this path); "
                "achieved = SURVEY 8(d) algorithmic flops per (pixel, Gaussian) pair x pairs / live kernel time; "
                "the pair count is upstream's walk (every list entry up to n_contrib), so frac > 1 means the "
                "kernel culled pairs upstream evaluates; executed-instruction evidence (smsp__issue_active, "
                "pipe_fma) is in profiles/",
    }
    kern = {"forward_ms": tf, "backward_ms": tb, "K": K, "P": P, "pairs_fwd": S, "pairs_bwd": S_bwd,
            "per_kernel": table}
    return roof, kern


def oracle_iteration(wl, P=None):
    """One fwd+bwd of the CPU oracle on the workload; returns a closure (inputs prepared once)."""
    import numpy as np
    import oracle
    from eags_raster import scene
    sc = scene.make_scene(wl, P=P, seed=0)
    cam = scene.workload_camera(wl)
    a = {k: sc[k].numpy() for k in sc}
    H, W = cam.H, cam.W

    def it():
        o = oracle.forward(W, H, cam.tanfovx, cam.tanfovy, cam.viewmatrix.numpy(), cam.projmatrix.numpy(),
                           cam.campos.numpy(), a["means3D"], a["scales"], a["rotations"], a["opacities"], shs=a["shs"])
        dC = np.sign(o["color"] - a["gt_color"]).astype(np.float32) / (3 * H * W)
        dD = np.sign(o["depth"] - a["gt_depth"]).astype(np.float32) / (H * W)
        oracle.backward(o, dC, dD, None)
    return it, cam, sc["means3D"].shape[0]


def cpu_baseline(wl, budget_s=15.0):
    import oracle
    oracle.build()
    cores = os.cpu_count() or 1
    oracle.set_num_threads(cores)
    it, cam, P = oracle_iteration(wl)
    it()   # warm-up (page faults, OpenMP pool)
    n, t0 = 0, time.perf_counter()
    while True:
        it()
        n += 1
        el = time.perf_counter() - t0
        if el > budget_s or n >= 20:
            break
    return {"value": n / el, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port",
            "sample": f"{n} full fwd+bwd iterations of {wl} ({cam.W}x{cam.H}, {P} Gaussians) with the C oracle (OpenMP)"}


def run_reference(args):
    """The reference arm: the reference's CUDA kernels are un-vendored pip packages and it has no
    CPU implementation of this path, so this times the CPU oracle port on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    oracle.build()
    cores = os.cpu_count() or 1
    oracle.set_num_threads(cores)
    it, cam, P = oracle_iteration(args.workload)
    steps = min(args.steps, 10)
    for _ in range(min(max(args.warmup, 1), 2)):
        it()
    t0 = time.perf_counter()
    for _ in range(steps):
        it()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    v = 1e3 / ms
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": min(max(args.warmup, 1), 2), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {cam.W}x{cam.H}, {P} Gaussians, fwd + L1 colour/depth loss + bwd"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port",
                         "sample": f"{steps} full fwd+bwd iterations per run (one iteration per step)"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sync-mode", default=None, choices=["optimistic", "exact", "nosync"])
    ap.add_argument("--tma", type=int, default=None)
    ap.add_argument("--loss", default="fused", choices=["torch", "fused"],
                    help="L1 colour + depth loss from torch ops, or the fused loss kernels of eags_raster.losses")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl"],
                    help="N > 1: fused peer-memory gradient exchange (default) or one NCCL all-reduce")
    ap.add_argument("--launch", default="graph", choices=["graph", "eager"],
                    help="N = 1: replay the step from a CUDA graph (default) or issue it eagerly every step")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    from eags_raster import _lib, rasterize as rz
    if args.sync_mode:
        rz.set_sync_mode(args.sync_mode)
    if args.tma is not None:
        _lib.set_option("tma", args.tma)
    run_ours(args)


if __name__ == "__main__":
    main()
