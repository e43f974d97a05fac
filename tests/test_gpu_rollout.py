"""GPU (-m gpu): the device rollout (resolution_pde_b200/rollout.py: rpde_rollout_renorm / rpde_rollout_loss behind
perform_rollout_1d/2d, rollout_losses, RolloutEvaluator) against the fixtures generated from the unmodified reference
(tests/golden/rollout), against the CPU oracle at other sizes, and through size-independent properties at benchmark sizes."""
import ctypes
import glob
import os

import numpy as np
import pytest
import torch

from helpers import TOL, build_candidate, rel
from rollout_helpers import ROLLOUT_GOLD, Norm, load_rollout_fixture, oracle_forward

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(ROLLOUT_GOLD, "rollout_*.npz")))
ENGINES = [e for e in os.environ.get("RPDE_TEST_PRECISION", "fp32,3xtf32").split(",") if e]
# a T-step rollout feeds every step's rounding back through the model: the per-call tolerance 1e-5 compounds over the steps
ROLL_TOL = 5 * TOL


@pytest.fixture(autouse=True, params=ENGINES)
def _engine(request):
    import resolution_pde_b200 as R
    R.set_precision(request.param)
    yield
    R.set_precision("fp32")


def _candidate(name, fx):
    m, _ = build_candidate(name[len("rollout_"):], fx["params"], fx["hyper"])
    m.load_state_dict(fx["params"])
    return m.to(DEV).eval()


def _norms(fx):
    return (Norm(fx["xn"]) if fx["xn"] else None), (Norm(fx["yn"]) if fx["yn"] else None)


@pytest.mark.parametrize("graph", [False, True])
@pytest.mark.parametrize("name", CASES)
def test_evaluator_matches_reference_fixture(name, graph):
    from resolution_pde_b200.rollout import RolloutEvaluator
    fx = load_rollout_fixture(name)
    xn, yn = _norms(fx)
    ev = RolloutEvaluator(_candidate(name, fx), fx["rollout_steps"], xn, yn, use_graph=graph)
    traj = fx["traj"].to(DEV)
    for _ in range(2):                       # the second call is a pure replay of the captured graph
        loss, steps, preds = ev(traj)
    assert tuple(preds.shape) == tuple(fx["preds"].shape)
    assert rel(preds, fx["preds"]) < ROLL_TOL
    np.testing.assert_allclose(steps.cpu().numpy(), fx["step_losses"], rtol=ROLL_TOL)
    assert abs(loss.item() - fx["batch_loss"]) < ROLL_TOL * abs(fx["batch_loss"])


@pytest.mark.parametrize("name", CASES)
def test_perform_rollout_matches_reference_fixture(name):
    from resolution_pde_b200.rollout import SimpleNormalizer, perform_rollout_1d, perform_rollout_2d, rollout_losses
    fx = load_rollout_fixture(name)
    xn = SimpleNormalizer(*fx["xn"]) if fx["xn"] else None
    yn = SimpleNormalizer(*fx["yn"]) if fx["yn"] else None
    traj = fx["traj"].to(DEV)
    T = fx["preds"].shape[1]
    ic = traj[:, 0]
    ic = xn.encode(ic.unsqueeze(1)).squeeze(1) if xn is not None else ic
    fn = perform_rollout_1d if traj.dim() == 3 else perform_rollout_2d
    preds = fn(_candidate(name, fx), ic, T, x_normalizer=xn, y_normalizer=yn)
    assert rel(preds, fx["preds"]) < ROLL_TOL
    loss, steps = rollout_losses(preds, traj, yn)
    np.testing.assert_allclose(steps.cpu().numpy(), fx["step_losses"], rtol=ROLL_TOL)
    assert abs(loss.item() - fx["batch_loss"]) < ROLL_TOL * abs(fx["batch_loss"])


def test_normalizer_is_bit_exact_with_reference_arithmetic():
    from resolution_pde_b200.rollout import SimpleNormalizer
    n = SimpleNormalizer(0.31, 1.7)
    for shape in ((3, 1, 64), (2, 1, 33), (2, 1, 16, 20), (5, 7)):
        x = torch.randn(*shape, generator=torch.Generator().manual_seed(sum(shape)))
        assert torch.equal(n.encode(x.to(DEV)).cpu(), (x - n.mean) / (n.std + n.eps))
        assert torch.equal(n.decode(x.to(DEV)).cpu(), x * (n.std + n.eps) + n.mean)


@pytest.mark.parametrize("shape,T", [((4, 6, 96), 5), ((2, 4, 31), 3), ((3, 5, 24, 40), 4), ((1, 3, 17, 9), 2)])
def test_renorm_and_loss_kernels_against_oracle(shape, T):
    """the two rollout kernels alone, driven by a trivial device 'model' (the state itself): stored predictions, the
    re-normalised next state (bit-exact) and the per-step losses (fp32 summation order differs: 1e-6)"""
    from oracle import rollout_port as rp
    from resolution_pde_b200.rollout import RolloutEvaluator
    g = torch.Generator().manual_seed(shape[-1] + T)
    traj = 0.4 + 1.3 * torch.randn(*shape, generator=g)
    xn, yn = (0.31, 1.7, 1e-8), (-0.22, 2.1, 1e-8)

    class Ident(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.p = torch.nn.Parameter(torch.zeros(1))

        def forward(self, x):
            return x

    ev = RolloutEvaluator(Ident().to(DEV), T, Norm(xn), Norm(yn), use_graph=False)
    loss, steps, preds = ev(traj.to(DEV))
    rl, rs, rpred = rp.rollout_batch_loss(lambda x: x, traj, T, xn, yn)
    assert torch.equal(preds.cpu(), rpred)
    np.testing.assert_allclose(steps.cpu().numpy(), rs.numpy(), rtol=2e-6)
    assert abs(loss.item() - rl.item()) < 2e-6 * abs(rl.item())


def test_rollout_other_resolutions_against_oracle():
    """super-resolution: the fixture's FNO1d / FFNO2D weights rolled out at grids they were not generated on"""
    from oracle import rollout_port as rp
    from resolution_pde_b200.rollout import RolloutEvaluator
    for name, shape in (("rollout_fno1d_n64_norm.npz", (2, 5, 256)), ("rollout_ffno1d_n32_norm.npz", (2, 4, 128)),
                        ("rollout_ffno2d_16x12_norm.npz", (2, 4, 32, 48)), ("rollout_fno2d_16x20_norm.npz", (1, 3, 64, 64))):
        fx = load_rollout_fixture(name)
        traj = 0.31 + 1.7 * torch.randn(*shape, generator=torch.Generator().manual_seed(shape[-1]))
        T = shape[1] - 1
        fwd = oracle_forward(name, {k: v.double() if v.is_floating_point() else v.to(torch.complex128)
                                    for k, v in fx["params"].items()}, fx["hyper"])
        with torch.no_grad():
            rl, rs, rpred = rp.rollout_batch_loss(fwd, traj.double(), T, fx["xn"], fx["yn"])
        xn, yn = _norms(fx)
        loss, steps, preds = RolloutEvaluator(_candidate(name, fx), T, xn, yn)(traj.to(DEV))
        assert rel(preds, rpred) < ROLL_TOL, name
        np.testing.assert_allclose(steps.cpu().numpy(), rs.numpy(), rtol=ROLL_TOL)


def test_evaluate_rollout_mean_over_batches_and_short_batches():
    from resolution_pde_b200.rollout import RolloutEvaluator, evaluate_rollout
    fx = load_rollout_fixture("rollout_fno1d_n64_norm.npz")
    xn, yn = _norms(fx)
    m = _candidate("rollout_fno1d_n64_norm.npz", fx)
    traj = fx["traj"].to(DEV)
    batches = [traj, traj[:2].contiguous(), traj[:, :1].contiguous()]      # the last has no step to roll out: skipped
    got = evaluate_rollout(m, batches, fx["rollout_steps"], xn, yn)
    ev = RolloutEvaluator(m, fx["rollout_steps"], xn, yn, use_graph=False)
    want = (ev(batches[0])[0].item() + ev(batches[1])[0].item()) / 2
    assert abs(got - want) < 1e-6 * abs(want)
    with pytest.raises(ValueError):
        ev(batches[2])


def test_full_size_properties():
    """C5 sizes: zero-weight-free properties that need no oracle.  (1) With identity re-normalisation (x and y statistics
    equal) the rollout equals the un-normalised one up to rounding; (2) a rollout of T steps starts with the rollout of
    T-1 steps (causality of the stored trajectory); (3) losses of a perfect prediction are zero."""
    import resolution_pde_b200 as R
    from resolution_pde_b200.rollout import perform_rollout_1d, rollout_losses
    torch.manual_seed(0)
    m = R.FNO1d(1, 1, modes=16, width=64).to(DEV).eval()
    x = torch.randn(32, 1024, device=DEV)
    p4 = perform_rollout_1d(m, x, 4)
    p3 = perform_rollout_1d(m, x, 3)
    assert rel(p4[:, :3], p3) < 1e-6                            # (split-K reductions may reorder sums between runs)
    n = Norm((0.0, 1.0, 0.0))
    pn = perform_rollout_1d(m, x, 4, x_normalizer=n, y_normalizer=n)
    assert rel(pn, p4) < 1e-6                                    # *1, +0, -0, /1 are exact
    traj = torch.cat([x.unsqueeze(1), p4], dim=1)
    loss, steps = rollout_losses(p4, traj)
    assert loss.item() == 0.0 and steps.abs().max().item() == 0.0


def test_cabi_rejects_bad_arguments():
    from resolution_pde_b200 import _lib
    lib = _lib.load()
    x = torch.zeros(8, device=DEV)
    assert lib.rpde_rollout_renorm(None, None, x.data_ptr(), 1, 8, 8, 8, 1.0, 0.0, 0.0, 1.0, 1, None) == -1
    assert lib.rpde_rollout_renorm(x.data_ptr(), None, x.data_ptr(), 1, 8, 4, 8, 1.0, 0.0, 0.0, 1.0, 1, None) == -1
    assert lib.rpde_rollout_renorm(x.data_ptr(), None, x.data_ptr(), 1, 8, 8, 8, 1.0, 0.0, 0.0, 0.0, 1, None) == -1
    assert lib.rpde_rollout_loss(x.data_ptr(), x.data_ptr(), None, x.data_ptr(), x.data_ptr(), 1, 1, 8, 8, 8, 8, 8, 1.0,
                                 0.0, 0, None) == -1
    assert b"rpde_rollout_loss" in lib.rpde_last_error()
