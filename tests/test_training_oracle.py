"""CPU: the restatement of the reference's training loop (oracle/training_port.py over the functional model ports) against
the loss histories written by the unmodified reference `train()` (oracle/make_golden_training.py)."""
import glob
import os

import numpy as np
import pytest

from training_helpers import TRAIN_GOLD, load_training_fixture, oracle_hyper

CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(TRAIN_GOLD, "train_*.npz")))


def test_fixtures_present():
    assert len(CASES) == 3


@pytest.mark.parametrize("name", CASES)
def test_training_port_reproduces_reference_loss_curves(name):
    from oracle import training_port as tp
    fx = load_training_fixture(name)
    kind, hyper = oracle_hyper(name, fx["hyper"])
    hist, vhist, _ = tp.train(kind, fx["params"], list(zip(fx["xs"], fx["ys"])), [(fx["vx"], fx["vy"])], fx["lr"], fx["epochs"], **hyper)
    np.testing.assert_allclose(hist, fx["loss_history"], rtol=1e-4)
    np.testing.assert_allclose(vhist, fx["val_history"], rtol=1e-4)
    assert fx["loss_history"][-1] < fx["loss_history"][0]            # 20 steps really train
