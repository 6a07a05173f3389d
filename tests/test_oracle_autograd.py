"""The C oracle's analytic backward (incl. the closed-form camera-pose gradient)
against float64 PyTorch autograd of the independent dense restatement oracle/torch_ref.py."""
import numpy as np
import pytest
import torch

import oracle
from oracle import torch_ref
from eags_raster import scene
from helpers import rel_err


@pytest.mark.parametrize("degree,M,variant,wl", [(0, 1, 0, "tiny"), (0, 1, 1, "tiny"), (3, 16, 0, "tiny"),
                                                 (0, 4, 1, "ragged")])
def test_oracle_backward_matches_autograd(degree, M, variant, wl):
    sc = scene.make_scene(wl, seed=3, M=M, outside=0.0, behind=0.05, scale_range=(0.7, 4.0))
    cam = scene.workload_camera(wl, w2c=scene.rot_xy(3, -2))
    H, W = cam.H, cam.W
    bg = (0.1, 0.2, 0.3)
    f = oracle.forward(W, H, cam.tanfovx, cam.tanfovy, cam.viewmatrix.numpy(), cam.projmatrix.numpy(),
                       cam.campos.numpy(), sc["means3D"].numpy(), sc["scales"].numpy(),
                       sc["rotations"].numpy(), sc["opacities"].numpy(), shs=sc["shs"].numpy(),
                       degree=degree, bg=bg, variant=variant)
    leaf = {k: sc[k].double().clone().requires_grad_(True)
            for k in ["means3D", "scales", "rotations", "opacities", "shs"]}
    tau = torch.zeros(6, dtype=torch.float64, requires_grad=True)
    r = torch_ref.render(W, H, cam.tanfovx, cam.tanfovy, cam.w2c, cam.projmatrix_raw, cam.campos,
                         leaf["means3D"], leaf["scales"], leaf["rotations"], leaf["opacities"],
                         shs=leaf["shs"], degree=degree, bg=bg, tau=tau)
    third = r["alpha_sum"] if variant == 0 else r["opacity"]
    assert (r["radii"].numpy() == f["radii"]).all()
    ok = f["fragile"] == 0
    assert ok.mean() > 0.99
    assert np.abs(r["color"].detach().numpy() - f["color"])[:, ok].max() < 5e-6
    assert np.abs(r["depth"].detach().numpy() - f["depth"][0])[ok].max() < 2e-5
    assert np.abs(third.detach().numpy() - f["third"][0])[ok].max() < 5e-6
    g = torch.Generator().manual_seed(7)
    dC = torch.randn(3, H, W, dtype=torch.float64, generator=g)
    dD = torch.randn(H, W, dtype=torch.float64, generator=g)
    dA = torch.randn(H, W, dtype=torch.float64, generator=g)
    fr = torch.from_numpy(~ok)
    dC[:, fr] = 0
    dD[fr] = 0
    dA[fr] = 0
    ((r["color"] * dC).sum() + (r["depth"] * dD).sum() + (third * dA).sum()).backward()
    gr = oracle.backward(f, dC.numpy(), dD.numpy(), dA.numpy(), want_tau=True)
    tol = 1e-4   # fp32 oracle vs f64 autograd; Q2's +1e-7 is below this
    assert rel_err(gr["dL_dmeans3D"], leaf["means3D"].grad.numpy()) < tol
    assert rel_err(gr["dL_dscales"], leaf["scales"].grad.numpy()) < tol
    assert rel_err(gr["dL_drots"], leaf["rotations"].grad.numpy()) < tol
    assert rel_err(gr["dL_dopacity"], leaf["opacities"].grad.numpy().reshape(-1)) < tol
    assert rel_err(gr["dL_dshs"], leaf["shs"].grad.numpy()) < tol
    assert rel_err(gr["dL_dtau"], tau.grad.numpy()) < tol
