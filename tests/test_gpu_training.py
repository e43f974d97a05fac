"""GPU (-m gpu): 20 optimizer steps of the drop-in modules against the loss histories the UNMODIFIED reference `train()`
produced from the same initial state-dict and batches (tests/golden/training; SURVEY 8c item 3: loss curves agree to 1e-4
relative).  Two drivers: the reference's loop as it is (torch AdamW + CosineAnnealingLR, eager), and the graph-captured step
with the fused AdamW (training.GraphedTrainStep + optim.FusedAdamW, lr pushed by the same scheduler)."""
import glob
import os

import numpy as np
import pytest
import torch

from helpers import build_candidate
from training_helpers import TRAIN_GOLD, load_training_fixture

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(TRAIN_GOLD, "train_*.npz")))
ENGINES = [e for e in os.environ.get("RPDE_TEST_PRECISION", "fp32,3xtf32").split(",") if e]
CURVE_TOL = 1e-4


@pytest.fixture(autouse=True, params=ENGINES)
def _engine(request):
    import resolution_pde_b200 as R
    R.set_precision(request.param)
    yield
    R.set_precision("fp32")


def _model(name, fx):
    m, _ = build_candidate(name[len("train_"):], fx["params"], fx["hyper"])
    m.load_state_dict(fx["params"])
    return m.to(DEV)


def _validate(m, loss_fn, fx):
    m.eval()
    with torch.no_grad():
        v = loss_fn(m(fx["vx"].to(DEV)), fx["vy"].to(DEV)).item()
    m.train()
    return v


@pytest.mark.parametrize("name", CASES)
def test_reference_loop_with_torch_adamw(name):
    """train/training.py:19-88 line by line on the drop-in modules"""
    import resolution_pde_b200 as R
    fx = load_training_fixture(name)
    m = _model(name, fx)
    loss_fn = R.RelativeL2Loss(size_average=True)
    opt = torch.optim.AdamW(m.parameters(), lr=fx["lr"], weight_decay=1e-4)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=100, eta_min=1e-5)
    hist, vhist = [], []
    for _ in range(fx["epochs"]):
        tot = 0.0
        for x, y in zip(fx["xs"], fx["ys"]):
            opt.zero_grad()
            loss = loss_fn(m(x.to(DEV)), y.to(DEV))
            loss.backward()
            opt.step()
            tot += loss.item()
        hist.append(tot / len(fx["xs"]))
        vhist.append(_validate(m, loss_fn, fx))
        sch.step()
    np.testing.assert_allclose(hist, fx["loss_history"], rtol=CURVE_TOL)
    np.testing.assert_allclose(vhist, fx["val_history"], rtol=CURVE_TOL)


@pytest.mark.parametrize("name", CASES)
def test_graphed_step_with_fused_adamw(name):
    """the same 20 steps as graph replays: zero_grad + forward + loss + backward in one graph, AdamW in a second, losses
    kept on the device (one sync per epoch)"""
    import resolution_pde_b200 as R
    fx = load_training_fixture(name)
    m = _model(name, fx)
    loss_fn = R.RelativeL2Loss(size_average=True)
    opt = R.FusedAdamW(m.parameters(), lr=fx["lr"], weight_decay=1e-4)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=100, eta_min=1e-5)
    step = R.GraphedTrainStep(m, loss_fn, optimizer=opt)
    hist, vhist = [], []
    for _ in range(fx["epochs"]):
        tot = torch.zeros((), device=DEV)
        for x, y in zip(fx["xs"], fx["ys"]):
            tot += step(x.to(DEV), y.to(DEV))
        hist.append(tot.item() / len(fx["xs"]))
        vhist.append(_validate(m, loss_fn, fx))
        sch.step()
    np.testing.assert_allclose(hist, fx["loss_history"], rtol=CURVE_TOL)
    np.testing.assert_allclose(vhist, fx["val_history"], rtol=CURVE_TOL)
