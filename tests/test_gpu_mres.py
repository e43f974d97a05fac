"""GPU (-m gpu): a true mixed-resolution training run (BASELINE config 4 at test size) through the device path.

The fixture (tests/golden/training/mres_fno2d.npz) holds the loss histories the UNMODIFIED reference produced with its own
ResolutionGroupedDataLoader + train() (train/mres_training.py:75-131, train/training.py:19-88) on a 16^2 / 24^2 / 32^2 mix with
ragged tail batches (4,4,2 / 4,1 / 3 samples).  The drop-in FNO2d must reproduce them to 1e-4 relative
  (a) through the reference's eager loop over R.ResolutionGroupedBatches with torch's AdamW,
  (b) through GraphedTrainStep + FusedAdamW: one CUDA graph per distinct (resolution, batch size) pair -- five graphs here,
and the bounded graph cache must evict without changing the result."""
import os

import numpy as np
import pytest
import torch

from test_mres_oracle import load

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ENGINES = [e for e in os.environ.get("RPDE_TEST_PRECISION", "fp32,3xtf32").split(",") if e]
CURVE_TOL = 1e-4


def _setup():
    import resolution_pde_b200 as R
    from resolution_pde_b200.training import ResolutionGroupedBatches
    z, samples, val, params = load()
    m = R.FNO2d(1, 1, modes1=4, modes2=4, width=8)
    m.load_state_dict(params)
    m = m.to(DEV)
    bs = int(z["batch_size"])
    tb = ResolutionGroupedBatches(samples, bs, shuffle=False, device=DEV)
    vb = ResolutionGroupedBatches(val, bs, shuffle=False, device=DEV)
    return R, z, m, tb, vb


def _validate(m, loss_fn, vb):
    m.eval()
    with torch.no_grad():
        v = sum(loss_fn(m(x), y).item() for x, y in vb) / len(vb)
    m.train()
    return v


@pytest.mark.parametrize("engine", ENGINES)
def test_mixed_resolution_epochs_eager(engine):
    R, z, m, tb, vb = _setup()
    R.set_precision(engine)
    try:
        loss_fn = R.RelativeL2Loss(size_average=True)
        opt = torch.optim.AdamW(m.parameters(), lr=float(z["lr"]), weight_decay=1e-4)
        sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=100, eta_min=1e-5)
        hist, vhist = [], []
        for _ in range(int(z["epochs"])):
            tot = 0.0
            for x, y in tb:
                opt.zero_grad()
                loss = loss_fn(m(x), y)
                loss.backward()
                opt.step()
                tot += loss.item()
            hist.append(tot / len(tb))
            vhist.append(_validate(m, loss_fn, vb))
            sch.step()
    finally:
        R.set_precision("fp32")
    np.testing.assert_allclose(hist, z["loss_history"], rtol=CURVE_TOL)
    np.testing.assert_allclose(vhist, z["val_history"], rtol=CURVE_TOL)


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("max_graphs", [16, 2])
def test_mixed_resolution_epochs_graphed(engine, max_graphs):
    R, z, m, tb, vb = _setup()
    R.set_precision(engine)
    try:
        loss_fn = R.RelativeL2Loss(size_average=True)
        opt = R.FusedAdamW(m.parameters(), lr=float(z["lr"]), weight_decay=1e-4)
        sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=100, eta_min=1e-5)
        step = R.GraphedTrainStep(m, loss_fn, optimizer=opt, max_graphs=max_graphs)
        hist, vhist = [], []
        for _ in range(int(z["epochs"])):
            tot = torch.zeros((), device=DEV)
            for x, y in tb:
                tot += step(x, y)
            hist.append(tot.item() / len(tb))
            vhist.append(_validate(m, loss_fn, vb))
            sch.step()
        assert len(step._graphs) == min(max_graphs, 5)       # one graph per distinct (resolution, batch size) pair, bounded
    finally:
        R.set_precision("fp32")
    np.testing.assert_allclose(hist, z["loss_history"], rtol=CURVE_TOL)
    np.testing.assert_allclose(vhist, z["val_history"], rtol=CURVE_TOL)
