"""Fused owner-side optimiser step of sharded mapping (csrc/collective.cu peer_adam_kernel,
eags_raster.peer.PeerAdam): sum over ranks -> chain rule through the reference's activations
(gaussian_model.py:77-81) -> isotropic regulariser (losses.py:128-138) -> Adam (gaussian_model.py:341-356)
-> parameter all-gather, against torch autograd + torch.optim.Adam on the same numbers.

`world` ranks are emulated on one device (cross-rank waits disabled, ranks run one after the other);
the real multi-GPU run is tests/peer_2gpu_check.py under torchrun."""
import pytest
import torch
import torch.nn.functional as F

from eags_raster import scene
from helpers import Settings

pytestmark = pytest.mark.gpu
LR = dict(xyz=1e-4, scaling=5e-3, rotation=1e-3, opacity=5e-2)


def _torch_reference(raw, eps, iso_w):
    ref = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    opt = torch.optim.Adam([{"params": [ref[k]], "lr": LR[k]} for k in ("xyz", "scaling", "rotation", "opacity")],
                           lr=0.0, eps=eps)
    return ref, opt


def _iso(scaling):
    m = scaling.mean(dim=1, keepdim=True)
    return (scaling - m * torch.ones_like(scaling)).abs().mean()      # reference losses.py:128-138


@pytest.mark.parametrize("world,P", [(1, 1000), (2, 4001), (3, 777), (8, 30_000)])
@pytest.mark.parametrize("eps", [1e-15, 1e-8])
def test_step_equals_torch_adam_on_given_gradients(world, P, eps):
    from eags_raster.peer import PeerAdam, PeerGradExchange, owner_chunk
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(P)
    raw = {"xyz": torch.randn(P, 3, generator=g), "scaling": torch.randn(P, 3, generator=g) - 3.0,
           "rotation": torch.randn(P, 4, generator=g) * 1.5, "opacity": torch.randn(P, 1, generator=g) * 2.0}
    raw = {k: v.to(dev) for k, v in raw.items()}
    ex = PeerGradExchange(P, dev, emulate_world=world)
    adam = PeerAdam(ex, lr=LR, eps=eps, isotropic_weight=1.0)
    for r in range(world):
        adam.load(raw["xyz"], raw["scaling"], raw["rotation"], raw["opacity"], rank=r)
    ref, opt = _torch_reference(raw, eps, 1.0)
    scale = 1.0 / (2 * world)
    names = {"means3D": "xyz", "scales": "scaling", "rotations": "rotation", "opacities": "opacity"}
    chunk = owner_chunk(P, world)
    for step in range(4):
        # what every rank's preprocess_bwd would have scattered: slot q of every owner = rank q's rows
        grads = [{k: torch.randn(ref[n].shape, generator=g).to(dev) * (10.0 ** float(torch.randint(-6, 0, (1,), generator=g)))
                  for k, n in names.items()} for _ in range(world)]
        for owner in range(world):
            lo, hi = min(P, owner * chunk), min(P, (owner + 1) * chunk)
            for q in range(world):
                rv = ex.recv_views(owner, q)
                for k in names:
                    rv[k][lo:hi].copy_(grads[q][k][lo:hi])
        for r in range(world):
            adam.step(scale, rank=r)
        # reference: d/d(raw) of  sum_k <act_k, scale * sum_q g_q,k> + isotropic_loss(exp(scaling))
        opt.zero_grad(set_to_none=True)
        act = {"means3D": ref["xyz"], "scales": torch.exp(ref["scaling"]), "rotations": F.normalize(ref["rotation"]),
               "opacities": torch.sigmoid(ref["opacity"])}
        total = _iso(act["scales"])
        for k in names:
            summed = grads[0][k].clone()
            for q in range(1, world):
                summed = summed + grads[q][k]
            total = total + (act[k] * (summed * scale)).sum()
        total.backward()
        opt.step()
        torch.cuda.synchronize()
        for r in range(world):
            got = adam.params(r)
            for n in LR:
                # an Adam update is at most ~lr per step: compare the parameters to 1e-3 of one update
                err = float((got[n] - ref[n].detach()).abs().max())
                assert err <= 1e-3 * LR[n] + 1e-6 * float(ref[n].detach().abs().max()), (step, r, n, err)
                assert torch.equal(got[n], adam.params(0)[n]), (step, r, n)      # replicas hold the owner's bits
            assert adam.steps_taken(r) == step + 1
    assert all(ex.error_flag(r) == 0 for r in range(world))


@pytest.mark.parametrize("world,per_rank,P", [(2, 2, 3001), (1, 3, 2000)])
def test_mapping_steps_through_the_rasterizer(world, per_rank, P):
    """The whole sharded step: every rank renders its views with the activated tensors of its replica, the
    last backward is routed with the fused optimiser.  Reference: autograd through the activations, mean over
    all views + isotropic loss, torch.optim.Adam.  eps = 1e-8 and weights scaled so that sqrt(v) >> eps is
    not assumed: the comparison is to a fraction of one update, for 99.9 % of the elements (a gradient whose
    sign is decided by the atomics' summation order moves its parameter by up to 2 lr at eps << |g|)."""
    from gaussian_rasterizer import GaussianRasterizer
    from eags_raster.peer import PeerAdam, PeerGradExchange
    dev = torch.device("cuda")
    sc = scene.make_scene("ragged", P=P, seed=7)
    views = scene.keyframe_views("ragged", 8)[:world * per_rank]
    raw = {"xyz": sc["means3D"].to(dev), "scaling": sc["scales"].to(dev).log(),
           "rotation": sc["rotations"].to(dev) * 1.7,
           "opacity": torch.logit(sc["opacities"].to(dev).clamp(1e-4, 1 - 1e-4))}
    shs = sc["shs"].to(dev)
    g = torch.Generator(device="cpu").manual_seed(3)
    weights = []
    for v in views:
        st = Settings(v, dev)
        weights.append((st, torch.randn(3, st.image_height, st.image_width, generator=g).to(dev) / 100,
                        torch.randn(1, st.image_height, st.image_width, generator=g).to(dev) / 100))

    def loss_of(act, i):
        st, wc, wd = weights[i]
        color, depth, alpha, radii = GaussianRasterizer(raster_settings=st)(
            means3D=act["means3D"], means2D=torch.zeros_like(act["means3D"], requires_grad=True),
            opacities=act["opacities"], shs=shs, colors_precomp=None, scales=act["scales"],
            rotations=act["rotations"], cov3D_precomp=None)
        return (color * wc).sum() + (depth * wd).sum()

    eps = 1e-8
    ref, opt = _torch_reference(raw, eps, 1.0)
    ex = PeerGradExchange(P, dev, emulate_world=world)
    adam = PeerAdam(ex, lr=LR, eps=eps)
    for r in range(world):
        adam.load(raw["xyz"], raw["scaling"], raw["rotation"], raw["opacity"], rank=r)
    scale = 1.0 / len(views)
    for step in range(3):
        for r in range(world):
            act = adam.activated(r)
            with ex.route(scale=scale, rank=r, reduce=False, local_views=per_rank, optimizer=adam):
                for j in range(per_rank):
                    loss_of(act, r * per_rank + j).backward()
        for r in range(world):
            adam.step(scale, rank=r)
        opt.zero_grad(set_to_none=True)
        act = {"means3D": ref["xyz"], "scales": torch.exp(ref["scaling"]), "rotations": F.normalize(ref["rotation"]),
               "opacities": torch.sigmoid(ref["opacity"])}
        total = _iso(act["scales"])
        for i in range(len(views)):
            total = total + scale * loss_of(act, i)
        total.backward()
        opt.step()
        torch.cuda.synchronize()
        for n in LR:
            got, want = adam.params(0)[n], ref[n].detach()
            close = (got - want).abs() <= 2e-2 * LR[n] + 1e-6 * want.abs()
            assert float(close.float().mean()) >= 0.999, (step, n, float(close.float().mean()))
            assert float((got - want).abs().max()) <= 2.5 * LR[n] * (step + 1), (step, n)
            for r in range(1, world):
                assert torch.equal(adam.params(r)[n], got), (step, r, n)
    assert all(ex.error_flag(r) == 0 for r in range(world))


def test_fused_activations_match_torch():
    """eags_gaussian_activations (one kernel) against the reference's activations (gaussian_model.py:77-81:
    torch.exp, torch.nn.functional.normalize, torch.sigmoid), including a zero quaternion (normalize's eps)
    and saturated opacities.  Tolerance: 2 ulp-ish relative (expf vs torch's exp), exact zeros stay zeros."""
    from eags_raster.peer import PeerAdam, PeerGradExchange
    dev = torch.device("cuda")
    P = 10_007
    g = torch.Generator().manual_seed(5)
    raw = {"xyz": torch.randn(P, 3, generator=g), "scaling": torch.randn(P, 3, generator=g) * 2 - 3,
           "rotation": torch.randn(P, 4, generator=g) * 3, "opacity": torch.randn(P, 1, generator=g) * 6}
    raw["rotation"][7] = 0.0
    raw["opacity"][:4, 0] = torch.tensor([-100.0, 100.0, 0.0, -20.0])
    raw = {k: v.to(dev) for k, v in raw.items()}
    ex = PeerGradExchange(P, dev, emulate_world=1)
    adam = PeerAdam(ex, lr=LR, eps=1e-15)
    adam.load(raw["xyz"], raw["scaling"], raw["rotation"], raw["opacity"])
    act = adam.activated()
    assert act["means3D"].data_ptr() == adam.params()["xyz"].data_ptr()      # xyz has no activation: a view
    want = {"scales": torch.exp(raw["scaling"]), "rotations": F.normalize(raw["rotation"]),
            "opacities": torch.sigmoid(raw["opacity"])}
    for k, w in want.items():
        assert act[k].shape == w.shape and act[k].is_contiguous()
        assert float(((act[k] - w).abs() - 4e-7 * w.abs()).max()) <= 1e-38, k
    assert float(act["rotations"][7].abs().max()) == 0.0
    assert not any(t.requires_grad for t in act.values())


@pytest.mark.parametrize("world", [1, 3])
def test_prune_and_grow_carry_the_optimiser_state(world):
    """PeerAdam.resized against the reference's optimiser surgery (gaussian_model.py:501-583: ``_prune_optimizer``
    keeps exp_avg / exp_avg_sq of the kept rows, ``cat_tensors_to_optimizer`` appends zeros, the step count is
    kept) on torch.optim.Adam with the same gradients: 2 steps, prune, 2 steps, grow, 2 steps; the ownership
    ranges move with P, so the moments cross emulated ranks."""
    from eags_raster.peer import PeerAdam, PeerGradExchange
    dev = torch.device("cuda")
    P, eps = 1501, 1e-8
    g = torch.Generator().manual_seed(11)
    raw = {"xyz": torch.randn(P, 3, generator=g), "scaling": torch.randn(P, 3, generator=g) - 3,
           "rotation": torch.randn(P, 4, generator=g), "opacity": torch.randn(P, 1, generator=g)}
    raw = {k: v.to(dev) for k, v in raw.items()}
    order = ("xyz", "scaling", "rotation", "opacity")
    ref = {k: torch.nn.Parameter(v.clone()) for k, v in raw.items()}
    opt = torch.optim.Adam([{"params": [ref[k]], "lr": LR[k], "name": k} for k in order], lr=0.0, eps=eps)
    ex = PeerGradExchange(P, dev, emulate_world=world)
    adam = PeerAdam(ex, lr=LR, eps=eps, isotropic_weight=0.0)
    for r in range(world):
        adam.load(raw["xyz"], raw["scaling"], raw["rotation"], raw["opacity"], rank=r)

    def both_step(seed):
        """The same RAW-parameter gradient on both sides: PeerAdam sees activated-input gradients, so feed it
        d(raw)/... through the inverse chain: give torch the chain-ruled gradient of what PeerAdam receives."""
        n = adam.ex.P
        gg = torch.Generator().manual_seed(seed)
        gact = {"means3D": torch.randn(n, 3, generator=gg).to(dev), "scales": torch.randn(n, 3, generator=gg).to(dev),
                "rotations": torch.randn(n, 4, generator=gg).to(dev), "opacities": torch.randn(n, 1, generator=gg).to(dev)}
        for r in range(world):          # every emulated rank sends gact / world
            for src in range(world):
                rv = adam.ex.recv_views(r, src)
                for k in gact:
                    rv[k].copy_(gact[k] / world)
        for r in range(world):
            adam.step(1.0, rank=r)
        opt.zero_grad(set_to_none=True)
        act = {"means3D": ref["xyz"], "scales": torch.exp(ref["scaling"]), "rotations": F.normalize(ref["rotation"]),
               "opacities": torch.sigmoid(ref["opacity"])}
        sum((act[k] * gact[k]).sum() for k in gact).backward()
        opt.step()

    def compare(tag):
        torch.cuda.synchronize()
        for n_ in order:
            got, want = adam.params(0)[n_], ref[n_].detach()
            assert got.shape == want.shape, (tag, n_)
            close = (got - want).abs() <= 2e-2 * LR[n_] + 1e-6 * want.abs()
            assert float(close.float().mean()) >= 0.999, (tag, n_, float(close.float().mean()))
            for r in range(1, world):
                assert torch.equal(adam.params(r)[n_], got), (tag, r, n_)

    def torch_surgery(keep=None, new=None):
        for group in opt.param_groups:
            p_old = group["params"][0]
            st = opt.state.pop(p_old)
            data, ea, es = p_old.detach(), st["exp_avg"], st["exp_avg_sq"]
            if keep is not None:
                data, ea, es = data[keep], ea[keep], es[keep]
            if new is not None:
                ext = new[group["name"]]
                data = torch.cat([data, ext]); ea = torch.cat([ea, torch.zeros_like(ext)])
                es = torch.cat([es, torch.zeros_like(ext)])
            p_new = torch.nn.Parameter(data.clone())
            st["exp_avg"], st["exp_avg_sq"] = ea.clone(), es.clone()
            group["params"][0] = p_new
            opt.state[p_new] = st
            ref[group["name"]] = p_new

    both_step(1); both_step(2); compare("before")
    keep = (torch.rand(P, generator=g) > 0.37).to(dev)
    ckpt = adam.state_dict()
    adam = adam.resized(keep=keep)
    torch_surgery(keep=keep)
    assert adam.ex.P == int(keep.sum()) and all(adam.steps_taken(r) == 2 for r in range(world))
    with pytest.raises(ValueError):
        adam.load_state_dict(ckpt)                      # taken before the prune
    compare("pruned")
    both_step(3); both_step(4); compare("after prune")
    n_new = 333
    new = {"xyz": torch.randn(n_new, 3, generator=g).to(dev), "scaling": (torch.randn(n_new, 3, generator=g) - 3).to(dev),
           "rotation": torch.randn(n_new, 4, generator=g).to(dev), "opacity": torch.randn(n_new, 1, generator=g).to(dev)}
    adam = adam.resized(new=new)
    torch_surgery(new=new)
    assert adam.ex.P == int(keep.sum()) + n_new
    compare("grown")
    sd = adam.state_dict()
    both_step(5); both_step(6); compare("after grow")
    adam.load_state_dict(sd)                            # restores moments + step count of this size
    assert all(adam.steps_taken(r) == 4 for r in range(world))
    with pytest.raises(ValueError):
        adam.resized(keep=torch.zeros(adam.ex.P, dtype=torch.bool, device=dev))
    assert all(adam.ex.error_flag(r) == 0 for r in range(world))
