"""torchrun --nproc-per-node N tests/peer_2gpu_check.py : the fused peer exchange against NCCL on real GPUs.
Every rank renders its own view of the same submap through the public API; p.grad after a routed
backward must equal the NCCL all-reduce of the plain backward's gradients x 1/world (<= 1 ulp-level
differences allowed only through the summation order: NCCL's order is its own, ours is rank order)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200"), os.path.join(ROOT, "tests")]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from eags_raster import scene  # noqa: E402
from eags_raster.peer import PeerGradExchange  # noqa: E402
from gaussian_rasterizer import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
P = 60_001
sc = scene.make_scene("c1", P=P, seed=0)
cam = scene.keyframe_views("c1", 8)[rank % 8]
s = GaussianRasterizationSettings(
    image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
    bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam.viewmatrix.to(dev),
    projmatrix=cam.projmatrix.to(dev), sh_degree=0, campos=cam.campos.to(dev), prefiltered=False, debug=False)
keys = ("means3D", "scales", "rotations", "opacities")
d = {k: sc[k].to(dev) for k in keys + ("shs", "gt_color", "gt_depth")}
raster = GaussianRasterizer(raster_settings=s)
ex = PeerGradExchange(P, dev)


def step(routed):
    leaf = {k: d[k].detach().requires_grad_(True) for k in keys}
    m2d = torch.zeros_like(leaf["means3D"], requires_grad=True)
    color, depth, alpha, radii = raster(means3D=leaf["means3D"], means2D=m2d, opacities=leaf["opacities"], shs=d["shs"],
                                        colors_precomp=None, scales=leaf["scales"], rotations=leaf["rotations"],
                                        cov3D_precomp=None)
    loss = (color - d["gt_color"]).abs().mean() + (depth - d["gt_depth"]).abs().mean()
    if routed:
        with ex.route(scale=1.0 / world):
            loss.backward()
        return {k: leaf[k].grad.clone() for k in keys}
    loss.backward()
    out = {}
    for k in keys:
        g = leaf[k].grad.clone()
        dist.all_reduce(g)
        out[k] = g / world
    return out


for it in range(3):          # several epochs: flags are never reset
    a, b = step(True), step(False)
    torch.cuda.synchronize()
    for k in keys:
        err = float((a[k] - b[k]).abs().max())
        ref = float(b[k].abs().max())
        assert err <= 1e-6 * ref + 1e-12, (it, k, err, ref)
        # replicas must hold the SAME bits: compare with rank 0's copy
        c = a[k].clone()
        dist.broadcast(c, src=0)
        assert torch.equal(c, a[k]), (it, k, "replicas differ")
assert ex.error_flag() == 0

# the same step replayed from a CUDA graph: the barrier epochs live in device memory, so every replay
# synchronises with its peers like an eager call; parameters change in place between replays
from eags_raster.graph import GraphedIteration  # noqa: E402

static = {k: d[k].clone().requires_grad_(True) for k in keys}


def graph_fn():
    for t in static.values():
        t.grad = None
    m2d = torch.zeros_like(static["means3D"], requires_grad=True)
    color, depth, alpha, radii = raster(means3D=static["means3D"], means2D=m2d, opacities=static["opacities"],
                                        shs=d["shs"], colors_precomp=None, scales=static["scales"],
                                        rotations=static["rotations"], cov3D_precomp=None)
    loss = (color - d["gt_color"]).abs().mean() + (depth - d["gt_depth"]).abs().mean()
    with ex.route(scale=1.0 / world):
        loss.backward()
    return {k: static[k].grad for k in keys}


gi = GraphedIteration(graph_fn, warmup=1)
for it in range(4):
    with torch.no_grad():
        static["means3D"].add_(1e-3)         # identical update on every rank
    g = gi.replay()
    gi.check()
    d_backup = {k: d[k] for k in keys}
    for k in keys:
        d[k] = static[k].detach()
    b = step(False)                           # eager NCCL reference on the updated parameters
    for k in keys:
        d[k] = d_backup[k]
    for k in keys:
        err, ref = float((g[k] - b[k]).abs().max()), float(b[k].abs().max())
        assert err <= 1e-5 * ref + 1e-12, ("graph", it, k, err, ref)
        c = g[k].clone()
        dist.broadcast(c, src=0)
        assert torch.equal(c, g[k]), ("graph", it, k, "replicas differ")
assert ex.error_flag() == 0

# several views per rank per step through ShardedMapper: the peer exchange (xyz / scaling / rotation /
# opacity; the rank's first view stays local, the second carries the sum) + NCCL for the SH features,
# against the pure NCCL mapper on identical replicas.  One SGD step each; parameters must agree.
from eags_raster import sharded  # noqa: E402

all_views = [scene.keyframe_views("c1", 8)[v % 8] for v in range(2 * world)]     # two views per rank at any world size


def settings_of(c):
    return GaussianRasterizationSettings(
        image_height=c.H, image_width=c.W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, bg=torch.zeros(3, device=dev),
        scale_modifier=1.0, viewmatrix=c.viewmatrix.to(dev), projmatrix=c.projmatrix.to(dev), sh_degree=0,
        campos=c.campos.to(dev), prefiltered=False, debug=False)


rasters = [GaussianRasterizer(raster_settings=settings_of(c)) for c in all_views]


def make_params():
    return {"xyz": d["means3D"].clone().requires_grad_(True), "scaling": d["scales"].log().requires_grad_(True),
            "rotation": (d["rotations"] * 1.3).clone().requires_grad_(True),
            "opacity": torch.logit(d["opacities"].clamp(1e-4, 1 - 1e-4)).requires_grad_(True),
            "features": d["shs"].clone().requires_grad_(True)}


def make_loss(prm):
    def loss_fn(v):
        m2d = torch.zeros_like(prm["xyz"], requires_grad=True)
        color, depth, alpha, radii = rasters[v](
            means3D=prm["xyz"], means2D=m2d, opacities=torch.sigmoid(prm["opacity"]), shs=prm["features"],
            colors_precomp=None, scales=torch.exp(prm["scaling"]),
            rotations=torch.nn.functional.normalize(prm["rotation"]), cov3D_precomp=None)
        return (color - d["gt_color"]).abs().mean() + (depth - d["gt_depth"]).abs().mean()
    return loss_fn


pa, pb = make_params(), make_params()
oa = torch.optim.SGD(list(pa.values()), lr=1e-2)
ob = torch.optim.SGD(list(pb.values()), lr=1e-2)
ma = sharded.ShardedMapper(list(pa.values()), oa, exchange=ex,
                           exchanged=[pa["xyz"], pa["scaling"], pa["rotation"], pa["opacity"]])
mb = sharded.ShardedMapper(list(pb.values()), ob)
views = list(range(len(all_views)))
for it in range(2):
    ra = ma.step(views, make_loss(pa), reg_fn=lambda: 0.1 * torch.exp(pa["scaling"]).mean())
    rb = mb.step(views, make_loss(pb), reg_fn=lambda: 0.1 * torch.exp(pb["scaling"]).mean())
    assert ra["local_views"] == rb["local_views"] == 2
    torch.cuda.synchronize()
    for k in pa:
        ga, gb = pa[k].grad, pb[k].grad
        err, ref = float((ga - gb).abs().max()), float(gb.abs().max())
        assert err <= 2e-5 * ref + 1e-12, ("mapper grad", it, k, err, ref)
        c = pa[k].detach().clone()
        dist.broadcast(c, src=0)
        assert torch.equal(c, pa[k].detach()), ("mapper", it, k, "replicas differ")
    # keep both mappers on the same trajectory (their gradients differ by summation order only)
    with torch.no_grad():
        for k in pa:
            pb[k].copy_(pa[k])
assert ex.error_flag() == 0

# the fused owner-side optimiser (eags_peer_adam_step) through ShardedMapper on real ranks: xyz / scaling /
# rotation / opacity live in the exchange and are stepped inside it; the SH features go through NCCL + SGD.
# Reference: the pure-NCCL mapper with torch Adam on identical replicas (eps 1e-8, no isotropic term).
from eags_raster.peer import PeerAdam  # noqa: E402

LR = dict(xyz=1e-4, scaling=5e-3, rotation=1e-3, opacity=5e-2)
ex3 = PeerGradExchange(P, dev)
adam = PeerAdam(ex3, lr=LR, eps=1e-8, isotropic_weight=0.0)
pc = make_params()
adam.load(pc["xyz"].detach(), pc["scaling"].detach(), pc["rotation"].detach(), pc["opacity"].detach())
feat_c = pc["features"]


def fused_loss(v):
    act = adam.activated()
    m2d = torch.zeros_like(act["means3D"], requires_grad=True)
    color, depth, alpha, radii = rasters[v](means3D=act["means3D"], means2D=m2d, opacities=act["opacities"], shs=feat_c,
                                            colors_precomp=None, scales=act["scales"], rotations=act["rotations"],
                                            cov3D_precomp=None)
    return (color - d["gt_color"]).abs().mean() + (depth - d["gt_depth"]).abs().mean()


pd_ = make_params()
od = torch.optim.Adam([{"params": [pd_[k]], "lr": LR[k]} for k in LR] + [{"params": [pd_["features"]], "lr": 0.0}], eps=1e-8)
mc = sharded.ShardedMapper([feat_c], torch.optim.SGD([feat_c], lr=0.0), exchange=ex3, fused_optimizer=adam, check_every=1)
md = sharded.ShardedMapper(list(pd_.values()), od)
for it in range(3):
    mc.step(views, fused_loss)
    md.step(views, make_loss(pd_))
    torch.cuda.synchronize()
    for k, lr in LR.items():
        a_, b_ = adam.params()[k], pd_[k].detach()
        close = ((a_ - b_).abs() <= 2e-2 * lr + 1e-6 * b_.abs()).float().mean()
        assert float(close) >= 0.999, ("fused adam", it, k, float(close))
        c = a_.clone()
        dist.broadcast(c, src=0)
        assert torch.equal(c, a_), ("fused adam", it, k, "replicas differ")
    with torch.no_grad():      # same trajectory for the next step
        for k in LR:
            pd_[k].copy_(adam.params()[k])
assert ex3.error_flag() == 0 and adam.steps_taken() == 3

# prune + grow on real ranks (ShardedMapper.resize -> PeerAdam.resized: one all-gather of the moments, a new
# symmetric-memory exchange): the kept Gaussians keep parameters and Adam moments, the new ones start at zero
# moments (gaussian_model.py:501-583).  Reference: the same surgery on the torch optimiser of the NCCL mapper.
gk = torch.Generator().manual_seed(77)
keep = (torch.rand(P, generator=gk) > 0.3).to(dev)
n_new = 1234
new = {"xyz": (torch.rand(n_new, 3, generator=gk) - 0.5).to(dev) * 2.0, "scaling": (torch.randn(n_new, 3, generator=gk) * 0.3 - 3.5).to(dev),
       "rotation": torch.randn(n_new, 4, generator=gk).to(dev), "opacity": torch.randn(n_new, 1, generator=gk).to(dev)}
new["xyz"][:, 2] += 3.0
new_feat = (torch.rand(n_new, feat_c.shape[1], 3, generator=gk) - 0.5).to(dev)
ex3, adam = mc.resize(keep=keep, new=new)
feat_c = torch.cat([feat_c.detach()[keep], new_feat]).requires_grad_(True)
mc.rebuild([feat_c], torch.optim.SGD([feat_c], lr=0.0))
names_ = ("xyz", "scaling", "rotation", "opacity")
for group, k in zip(od.param_groups[:4], names_):
    p_old = group["params"][0]
    st = od.state.pop(p_old)
    data = torch.cat([p_old.detach()[keep], new[k]])
    for key in ("exp_avg", "exp_avg_sq"):
        st[key] = torch.cat([st[key][keep], torch.zeros_like(new[k])])
    p_new = torch.nn.Parameter(data.clone())
    group["params"][0] = p_new
    od.state[p_new] = st
    pd_[k] = p_new
fg = od.param_groups[4]
od.state.pop(fg["params"][0], None)
pd_["features"] = torch.nn.Parameter(feat_c.detach().clone())
fg["params"][0] = pd_["features"]
md.rebuild(list(pd_.values()), od)
P2 = int(keep.sum()) + n_new
assert adam.ex.P == P2 and adam.steps_taken() == 3
for it in range(2):
    mc.step(views, fused_loss)
    md.step(views, make_loss(pd_))
    torch.cuda.synchronize()
    for k, lr in LR.items():
        a_, b_ = adam.params()[k], pd_[k].detach()
        close = ((a_ - b_).abs() <= 2e-2 * lr + 1e-6 * b_.abs()).float().mean()
        assert float(close) >= 0.999, ("fused adam after resize", it, k, float(close))
        c = a_.clone()
        dist.broadcast(c, src=0)
        assert torch.equal(c, a_), ("fused adam after resize", it, k, "replicas differ")
    with torch.no_grad():
        for k in LR:
            pd_[k].copy_(adam.params()[k])
assert ex3.error_flag() == 0 and adam.steps_taken() == 5
dist.barrier()
if rank == 0:
    print(f"prune + grow ok: {P} -> {P2} Gaussians on {world} ranks, optimiser state carried over, 2 more steps vs torch Adam")
    print(f"fused peer Adam ok: 3 steps of {len(views)} views on {world} ranks vs NCCL + torch Adam, replicas bit-identical")
    print(f"sharded mapper ok: {len(views)} views on {world} ranks (2 per rank), peer exchange + NCCL features vs NCCL only")
    print(f"peer exchange ok: world {world}, P {P}, 3 eager steps + 4 graph replays, max rel diff vs NCCL <= 1e-5, replicas bit-identical")
dist.destroy_process_group()
