"""CPU: the restatement of the reference's autoregressive rollout (oracle/rollout_port.py, driving the model ports of
oracle/reference_port.py) against the fixtures generated from the unmodified reference perform_rollout_1d, models and
RelativeL2Loss (oracle/make_golden_rollout.py)."""
import glob
import os

import numpy as np
import pytest
import torch

from rollout_helpers import ROLLOUT_GOLD, load_rollout_fixture, oracle_forward

CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(ROLLOUT_GOLD, "rollout_*.npz")))


def test_fixtures_present():
    assert len(CASES) == 6


@pytest.mark.parametrize("name", CASES)
def test_rollout_port_matches_reference(name):
    from oracle import rollout_port as rp
    fx = load_rollout_fixture(name)
    fwd = oracle_forward(name, fx["params"], fx["hyper"])
    with torch.no_grad():
        loss, steps, preds = rp.rollout_batch_loss(fwd, fx["traj"], fx["rollout_steps"], fx["xn"], fx["yn"])
    assert preds.shape == fx["preds"].shape
    assert ((preds - fx["preds"]).norm() / fx["preds"].norm()).item() < 2e-6
    np.testing.assert_allclose(steps.numpy(), fx["step_losses"], rtol=2e-6)
    assert abs(loss.item() - fx["batch_loss"]) < 2e-6 * abs(fx["batch_loss"])


def test_rollout_port_clamps_and_skips():
    from oracle import rollout_port as rp
    ident = lambda x: x
    traj = torch.arange(2 * 3 * 4, dtype=torch.float32).reshape(2, 3, 4) + 1
    loss, steps, preds = rp.rollout_batch_loss(ident, traj, 10)
    assert preds.shape == (2, 2, 4) and steps.shape == (2,)
    assert torch.equal(preds[:, 0], traj[:, 0]) and torch.equal(preds[:, 1], traj[:, 0])
    with pytest.raises(ValueError):
        rp.rollout_batch_loss(ident, traj[:, :1], 3)
    assert rp.evaluate_rollout(ident, [traj, traj[:, :1]], 10) == pytest.approx(loss.item())


def test_renormalisation_is_decode_then_encode():
    from oracle import rollout_port as rp
    xn, yn = (0.31, 1.7, 1e-8), (-0.22, 2.1, 1e-8)
    x = torch.randn(3, 16, generator=torch.Generator().manual_seed(0))
    p = rp.perform_rollout(lambda s: 0.5 * s, x, 2, xn, yn)
    s1 = ((0.5 * x.unsqueeze(1)).squeeze(1) * (2.1 + 1e-8) + (-0.22) - 0.31) / (1.7 + 1e-8)
    assert torch.equal(p[:, 1], 0.5 * s1)
