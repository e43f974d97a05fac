"""Fused multi-tensor AdamW (rpde_adamw_step / FusedAdamW) against torch.optim.AdamW, the optimizer the reference
builds (main_1d.py:144, main_2d.py:173).  Tolerance 2e-6 relative L2 on parameters and moments after several steps."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel(a, b):
    a, b = torch.view_as_real(a) if a.is_complex() else a, torch.view_as_real(b) if b.is_complex() else b
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def _params(seed):
    torch.manual_seed(seed)
    shapes = [(1,), (3,), (1023,), (1024,), (1025,), (37, 129), (64, 64, 3)]
    ps = [torch.randn(s, device=DEV) for s in shapes]
    ps.append(torch.randn(7, 5, 9, dtype=torch.cfloat, device=DEV))          # FNO spectral weights are complex64
    ps.append(torch.randn(2, 3, dtype=torch.cfloat, device=DEV))
    return [torch.nn.Parameter(p) for p in ps]


@pytest.mark.parametrize("wd,betas", [(1e-2, (0.9, 0.999)), (0.0, (0.8, 0.95)), (1e-4, (0.9, 0.999))])
def test_fused_adamw_matches_torch(wd, betas):
    import resolution_pde_b200 as R
    pa, pb = _params(1), _params(1)
    oa = R.FusedAdamW(pa, lr=3e-3, betas=betas, eps=1e-8, weight_decay=wd)
    ob = torch.optim.AdamW(pb, lr=3e-3, betas=betas, eps=1e-8, weight_decay=wd)
    for it in range(7):
        torch.manual_seed(100 + it)
        for a, b in zip(pa, pb):
            g = torch.randn_like(a) * (10.0 ** (it % 3 - 1))
            a.grad, b.grad = g.clone(), g.clone()                             # fresh tensors: the device table is rebuilt
        if it == 3:                                                            # what a scheduler does
            oa.param_groups[0]["lr"] = ob.param_groups[0]["lr"] = 1e-3
        oa.step()
        ob.step()
    for a, b in zip(pa, pb):
        assert rel(a, b) < 2e-6
        assert rel(oa.state[a]["exp_avg"], ob.state[b]["exp_avg"]) < 2e-6
        assert rel(oa.state[a]["exp_avg_sq"], ob.state[b]["exp_avg_sq"]) < 2e-6
    assert float(oa.state_dict()["state"][0]["step"]) == 7.0


def test_state_dict_round_trip_with_torch_adamw():
    import resolution_pde_b200 as R
    pa, pb = _params(2), _params(2)
    oa = R.FusedAdamW(pa, lr=2e-3, weight_decay=1e-2)
    for it in range(3):
        for a in pa:
            a.grad = torch.randn_like(a)
        oa.step()
    with torch.no_grad():
        for a, b in zip(pa, pb):
            b.copy_(a)
    ob = torch.optim.AdamW(pb, lr=2e-3, weight_decay=1e-2)
    ob.load_state_dict(copy.deepcopy(oa.state_dict()))                         # fused -> torch
    oc = R.FusedAdamW(_params(2), lr=2e-3, weight_decay=1e-2)
    pc = oc.param_groups[0]["params"]
    with torch.no_grad():
        for a, c in zip(pa, pc):
            c.copy_(a)
    oc.load_state_dict(copy.deepcopy(ob.state_dict()))                         # torch -> fused
    for it in range(2):
        torch.manual_seed(50 + it)
        for a, b, c in zip(pa, pb, pc):
            g = torch.randn_like(a)
            a.grad, b.grad, c.grad = g.clone(), g.clone(), g.clone()
        oa.step(); ob.step(); oc.step()
    for a, b, c in zip(pa, pb, pc):
        assert rel(a, b) < 2e-6 and rel(c, b) < 2e-6


def test_graphed_step_with_optimizer_matches_eager_torch_adamw():
    import resolution_pde_b200 as R
    R.set_precision("fp32")
    torch.manual_seed(3)
    # two input channels: with one, weight_v of the weight-normed in_proj has an analytically zero gradient and Adam
    # turns its rounding noise into +-lr steps, which no two implementations share
    ma = R.FFNO1D(2, 1, width=32, n_layers=2, n_modes=8, factor=2, ff_weight_norm=True, n_ff_layers=2, layer_norm=True,
                  dropout=0.0).to(DEV)
    mb = copy.deepcopy(ma)
    loss_fn = R.RelativeL2Loss()
    oa = R.FusedAdamW(ma.parameters(), lr=1e-3, weight_decay=1e-2)
    ob = torch.optim.AdamW(mb.parameters(), lr=1e-3, weight_decay=1e-2)
    step = R.GraphedTrainStep(ma, loss_fn, optimizer=oa)
    for it in range(4):
        torch.manual_seed(10 + it)
        x, y = torch.randn(4, 2, 64, device=DEV), torch.randn(4, 1, 64, device=DEV)
        la = step(x, y)
        mb.zero_grad(set_to_none=True)
        lb = loss_fn(mb(x), y)
        lb.backward()
        ob.step()
        assert abs(la.item() - lb.item()) < 1e-5 * max(1.0, abs(lb.item()))
    for (k, a), (_, b) in zip(ma.named_parameters(), mb.named_parameters()):
        assert rel(a, b) < 1e-5, k
