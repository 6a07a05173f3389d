"""Parallel loop closure (BASELINE.json configs[4], SURVEY 8(e) c5 + 8(f) N-4): 16 independent ~300 K-Gaussian
submaps placed over the GPUs of one box, no collective on the data path.

Per submap, as the reference's registration does (src/gsr/solver.py:16-98 viewpoint_localizer): for each of
its keyframe views, `--iters` pose-gradient iterations through variant B of the operator
(`diff_gaussian_rasterization`: theta / rho gradients, opacity image, n_touched), Adam on the two pose deltas
(lr 3e-3 / 1e-3), then the camera update `w2c <- SE3_exp(tau) w2c` with the deltas zeroed
(pose_utils.py:76-93) -- here one fused kernel on the device (eags_raster.tracking.update_pose_se3) so the
whole iteration replays from a CUDA graph.  Loss: opacity-weighted L1 colour + masked L1 depth
(src/gsr/loss.py:66-95).

Hand-off (N-4): the submaps are produced on rank 0 (the mapping process) and delivered to the rank that
registers them with eags_raster.handoff.send_submap / recv_submap (one NCCL send per (dtype, device) block,
NVLink) -- timed against the reference's path, torch.save on the producer + torch.load(map_location=device)
on the consumer (src/entities/submap.py:20-37,109-124).  Inside ONE process (the reference's layout: mapping
thread on device 0, loop-closure thread on lc.device) the same dictionary goes through
SubmapMailbox.put(cuda:0 -> cuda:k); rank 0 times that too when the box shows it a second device.

usage: [python -m torch.distributed.run --nproc-per-node N] python profiles/bench_lc.py [--submaps 16] [--views 4] [--iters 100]
"""
import argparse
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402
from eags_raster import handoff, scene  # noqa: E402
from eags_raster.graph import GraphedIteration  # noqa: E402
from eags_raster.sharded import assign_submaps  # noqa: E402
from eags_raster.losses import registration_loss  # noqa: E402
from eags_raster.tracking import PoseAdamSE3, update_pose_se3  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--submaps", type=int, default=16)
ap.add_argument("--views", type=int, default=4)
ap.add_argument("--iters", type=int, default=100)
ap.add_argument("--P", type=int, default=300_000)
ap.add_argument("--launch", default="graph", choices=["graph", "eager"])
ap.add_argument("--host-side", default="fused", choices=["fused", "torch"],
                help="loss + optimiser + camera update: the fused kernels (registration_loss, PoseAdamSE3) or torch ops + "
                     "torch.optim.Adam + update_pose_se3")
ap.add_argument("--json", default=None)
a = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def make_submap(sid):
    """The dictionary Submap.from_model / save produce (submap.py:72-124): gaussian_params + Tmc on the mapping
    GPU, T_prev_m (float64) and kf_id (int32) on the CPU."""
    sc = scene.make_scene("c5", P=a.P, seed=sid)
    gp = {"xyz": sc["means3D"].to(dev), "features_dc": sc["shs"].to(dev), "features_rest": torch.zeros(a.P, 0, 3, device=dev),
          "scaling": sc["scales"].log().to(dev), "rotation": sc["rotations"].to(dev),
          "opacity": torch.logit(sc["opacities"].clamp(1e-4, 1 - 1e-4)).to(dev)}
    return {"id": sid, "gaussian_params": gp, "Tmc": torch.eye(4, dtype=torch.float64, device=dev)[None].repeat(a.views, 1, 1),
            "kf_id": torch.arange(sid * a.views, (sid + 1) * a.views).int(), "T_prev_m": torch.eye(4, dtype=torch.float64)}


# ---- phase A: hand-off of every submap from rank 0 to the rank that registers it -------------------------------
mine = assign_submaps(a.submaps, rank, world)
owned, t_hand, t_disk_save, t_disk_load, t_box = {}, [], [], [], []
tmp = tempfile.mkdtemp(prefix="eags_lc_") if rank == 0 else None
if world > 1:
    box_ = [tmp]
    dist.broadcast_object_list(box_, src=0)
    tmp = box_[0]
for sid in range(a.submaps):
    dst = sid % world
    if rank == 0:
        sm = make_submap(sid)
        torch.cuda.synchronize()
        if dst == 0:
            owned[sid] = sm
        else:
            ack = torch.zeros(1, device=dev)
            t0 = time.perf_counter()
            handoff.send_submap(sm, dst=dst)
            dist.recv(ack, src=dst)                   # the consumer holds the complete submap on its device
            torch.cuda.synchronize()
            t_hand.append(time.perf_counter() - t0)
            # the reference's path: a checkpoint file written by the producer, read by the consumer
            t0 = time.perf_counter()
            handoff.to_ckpt(sm, tmp)
            t_disk_save.append(time.perf_counter() - t0)
            dist.send(torch.ones(1, device=dev), dst=dst)
            dist.recv(ack, src=dst)
        if sid == 0 and torch.cuda.device_count() > 1:
            # one process, two devices (the reference's thread layout): SubmapMailbox.put(cuda:0 -> cuda:k)
            other = torch.device("cuda", (local + 1) % torch.cuda.device_count())
            box = handoff.SubmapMailbox()
            for rep in range(3):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                box.put(sm, other)
                got = box.get(sid, pop=True)
                t_box.append(time.perf_counter() - t0)
            assert all(t.device == other for _, t in handoff._walk(got))
            assert torch.equal(got["gaussian_params"]["xyz"].to(dev), sm["gaussian_params"]["xyz"])
            del got, box
    elif rank == dst:
        got = handoff.recv_submap(0, dev)
        torch.cuda.synchronize()
        dist.send(torch.ones(1, device=dev), dst=0)
        owned[sid] = got
        dist.recv(torch.zeros(1, device=dev), src=0)  # the checkpoint file is complete ...
        torch.cuda.synchronize()                      # ... once the (stream-ordered) NCCL receive has really landed
        t0 = time.perf_counter()
        back = torch.load(os.path.join(tmp, f"{str(sid).zfill(6)}.ckpt"), map_location=dev, weights_only=False)
        torch.cuda.synchronize()
        t_disk_load.append(time.perf_counter() - t0)
        ok = all(torch.equal(got["gaussian_params"][k], back["gaussian_params"][k]) for k in back["gaussian_params"])
        assert ok and torch.equal(got["Tmc"], back["Tmc"]) and torch.equal(got["kf_id"].cpu(), back["kf_id"].cpu())
        dist.send(torch.ones(1, device=dev), dst=0)
        del back

# ---- phase B: registration of the owned submaps ----------------------------------------------------------------
cams = scene.keyframe_views("c5", max(a.views, 1))
H, W = cams[0].H, cams[0].W
total_ms, total_it, final_norms = 0.0, 0, []
for sid in mine:
    sm = owned.pop(sid)
    gp = sm["gaussian_params"]
    d = {"means3D": gp["xyz"], "scales": torch.exp(gp["scaling"]), "rotations": torch.nn.functional.normalize(gp["rotation"]),
         "opacities": torch.sigmoid(gp["opacity"]), "shs": torch.cat((gp["features_dc"], gp["features_rest"]), dim=1)}
    gsc = scene.make_scene("c5", P=16, seed=1000 + sid)           # only the synthetic keyframe images are used
    gt_c, gt_d = gsc["gt_color"].to(dev), gsc["gt_depth"].to(dev)
    for v in range(a.views):
        cam = cams[v]
        view = cam.viewmatrix.to(dev).contiguous().clone()
        proj, proj_raw, campos = cam.projmatrix.to(dev).contiguous().clone(), cam.projmatrix_raw.to(dev).contiguous(), cam.campos.to(dev).clone()
        s = GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=torch.zeros(3, device=dev),
            scale_modifier=1.0, viewmatrix=view, projmatrix=proj, projmatrix_raw=proj_raw, sh_degree=0, campos=campos,
            prefiltered=False, debug=False)
        raster = GaussianRasterizer(raster_settings=s)
        theta = torch.zeros(3, device=dev, requires_grad=True)
        rho = torch.zeros(3, device=dev, requires_grad=True)
        fused = a.host_side == "fused"
        opt = PoseAdamSE3(dev, base_lr=1e-3) if fused else \
            torch.optim.Adam([{"params": [theta], "lr": 3e-3}, {"params": [rho], "lr": 1e-3}], capturable=True)
        tau_norm = torch.zeros(1, device=dev)
        rgb_mask_b, depth_mask_b = gt_c.sum(dim=0) > 0.01, gt_d[0] > 0.01
        rgb_mask, depth_mask = rgb_mask_b[None].float(), depth_mask_b[None].float()

        def iteration():
            if theta.grad is not None:
                theta.grad.zero_()
                rho.grad.zero_()
            m2d = torch.zeros_like(d["means3D"])
            color, radii, depth, opacity, n_touched = raster(
                means3D=d["means3D"], means2D=m2d, opacities=d["opacities"], shs=d["shs"], scales=d["scales"],
                rotations=d["rotations"], theta=theta, rho=rho)
            # get_loss_tracking_rgbd (src/gsr/loss.py:66-95), alpha = 0.95
            if fused:
                loss = registration_loss(color, depth, opacity, gt_c, gt_d, rgb_mask_b, depth_mask_b)
            else:
                l1_rgb = (opacity * (color * rgb_mask - gt_c * rgb_mask).abs()).mean()
                dm = depth_mask * (opacity > 0.95).float()
                loss = 0.95 * l1_rgb + 0.05 * (depth * dm - gt_d * dm).abs().mean()
            loss.backward()
            with torch.no_grad():
                if fused:
                    opt.step(theta, rho, view, proj_raw, proj, campos, tau_norm)
                else:
                    opt.step()
                    update_pose_se3(theta, rho, view, proj_raw, proj, campos, tau_norm)
            return loss.detach()

        for _ in range(2):
            iteration()
        run = iteration
        if a.launch == "graph":
            gi = GraphedIteration(iteration, warmup=1)
            run = gi.replay
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            run()
        e1.record()
        torch.cuda.synchronize()
        if a.launch == "graph":
            gi.check()
        total_ms += e0.elapsed_time(e1)
        total_it += a.iters
        final_norms.append(float(tau_norm))
        assert torch.isfinite(view).all() and torch.isfinite(proj).all()
    del sm, gp, d
t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    ms = float(t)
    med = lambda x: (sorted(x)[len(x) // 2] * 1e3) if x else None     # noqa: E731
    ld = None
    if world > 1:
        ldt = torch.zeros(1, device=dev, dtype=torch.float64)
        dist.recv(ldt, src=1)
        ld = float(ldt)
    bytes_sm = a.P * (3 + 3 + 3 + 4 + 1) * 4
    res = {"workload": f"c5: {a.submaps} independent submaps x {a.P} Gaussians, 640x480, {a.views} views x {a.iters} "
                       f"variant-B pose iterations each (Adam + SE3 update per iteration)",
           "n_gpus": world, "launch": a.launch, "host_side": a.host_side, "value": a.submaps / ms * 1e3, "unit": "submaps/s",
           "iterations_per_s": a.submaps * a.views * a.iters / ms * 1e3, "ms_per_iteration_per_gpu": ms / max(total_it, 1),
           "collective": "none (replicas only)",
           "handoff": {"submap_bytes": bytes_sm,
                       "send_recv_submap_ms_median": med(t_hand), "send_recv_submap_ms_min": (min(t_hand) * 1e3) if t_hand else None,
                       "send_recv_submap_ms_all": [round(x * 1e3, 3) for x in t_hand],
                       "disk_torch_save_ms_median": med(t_disk_save), "disk_torch_load_ms_median": ld,
                       "mailbox_put_other_device_ms": [x * 1e3 for x in t_box] or None,
                       "note": "send/recv: rank 0 -> registering rank over NCCL P2P incl. the pickled header and the "
                               "consumer's acknowledgement (the FIRST send to each peer also creates the NCCL P2P channel, "
                               "hundreds of ms: see _all); disk: the reference's torch.save + torch.load(map_location) of the "
                               "same dictionary through a temp directory; mailbox: one process, cuda:0 -> another device"}}
    print(json.dumps(res))
    if a.json:
        with open(a.json, "w") as f:
            json.dump(res, f, indent=1)
elif rank == 1:
    x = sorted(t_disk_load)[len(t_disk_load) // 2] * 1e3 if t_disk_load else 0.0
    dist.send(torch.tensor([x], device=dev, dtype=torch.float64), dst=0)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
