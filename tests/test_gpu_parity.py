"""GPU parity tests (run on the B200 box with -m gpu): the CUDA path, called through the drop-in modules and the
C ABI, against (1) the reference-generated golden fixtures, (2) the CPU oracle evaluated in fp64 on seeded inputs
across resolutions, (3) size-independent properties at full benchmark sizes."""
import glob
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import GOLD, TOL, assert_grads_close, build_candidate, load_fixture, rel

pytestmark = pytest.mark.gpu
CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*.npz")) if "relative_l2" not in p)
DEV = "cuda:0"


# Every parity test runs on both fp32-accurate engines: the CUDA-core FMA engine ("fp32") and the tcgen05 3xTF32
# engine ("3xtf32"); the tolerance is the same 1e-5 (BASELINE.json north_star).  RPDE_TEST_PRECISION restricts the set.
ENGINES = [e for e in os.environ.get("RPDE_TEST_PRECISION", "fp32,3xtf32").split(",") if e]


@pytest.fixture(autouse=True, params=ENGINES)
def _engine(request):
    import resolution_pde_b200 as R
    R.set_precision(request.param)
    yield
    R.set_precision("fp32")


def _step(m, x, y, out_index=None):
    import resolution_pde_b200 as R
    x = x.clone().requires_grad_(True)
    out = m(x)
    if out_index is not None:
        out = out[out_index]
    loss = R.RelativeL2Loss()(out, y)
    loss.backward()
    return out.detach(), loss.detach(), x.grad, {k: (p.grad if p.grad is not None else torch.zeros_like(p))
                                                 for k, p in m.named_parameters()}


@pytest.mark.parametrize("name", CASES)
def test_golden_fixture(name):
    z, params, grads, hyper = load_fixture(name)
    m, oi = build_candidate(name, params, hyper)
    m.load_state_dict(params)
    m = m.to(DEV)
    out, loss, gx, g = _step(m, torch.from_numpy(z["x"]).to(DEV), torch.from_numpy(z["y"]).to(DEV), oi)
    assert rel(out, z["out"]) < TOL
    assert abs(loss.item() - float(z["loss"])) < TOL * max(1.0, abs(float(z["loss"])))
    assert rel(gx, z["grad_x"]) < TOL
    assert_grads_close(g, grads)


def test_relative_l2_fixture():
    import resolution_pde_b200 as R
    z = np.load(os.path.join(GOLD, "relative_l2.npz"))
    a, b = torch.from_numpy(z["x"]).to(DEV), torch.from_numpy(z["y"]).to(DEV)
    assert rel(R.RelativeL2Loss(True, True)(a, b), z["mean"]) < 1e-6
    assert rel(R.RelativeL2Loss(False, True)(a, b), z["sum"]) < 1e-6
    assert rel(R.RelativeL2Loss(True, False)(a, b), z["none"]) < 1e-6


# ---------------------------------------------------------------------------------------------------------
# oracle (fp64 on the CPU) vs CUDA path across resolutions
# ---------------------------------------------------------------------------------------------------------
def _oracle_step(kind, sd, x, y, **hyper):
    from oracle import reference_port as rp
    return rp.step_loss_and_grads(kind, sd, x, y, dtype=torch.float64, **hyper)


def _compare_model(kind, m, x, y, hyper, tol=TOL):
    sd = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    pred_o, loss_o, g_o = _oracle_step(kind, sd, x, y, **hyper)
    out, loss, gx, g = _step(m.to(DEV), x.to(DEV), y.to(DEV))
    assert rel(out, pred_o) < tol, rel(out, pred_o)
    assert abs(loss.item() - loss_o.item()) < tol * max(1.0, abs(loss_o.item()))
    assert_grads_close(g, {k: v for k, v in g_o.items()}, tol)


def _relu_margin(kind, m, x, y, hyper):
    """Smallest |pre-activation| / rms over every ReLU input of the fp64 oracle.  Below fp32 resolution the mask of that
    element is a coin toss in any fp32 implementation (the gradient is discontinuous there), so such a draw cannot
    be held to 1e-5; the resolution tests skip to the next seed instead."""
    import torch.nn.functional as F
    from oracle import reference_port as rp
    margins = []

    def rec(t, *a, **k):
        margins.append((t.abs().min() / t.pow(2).mean().sqrt()).item())
        return F.relu(t, *a, **k)

    sd = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    rp.step_loss_and_grads(kind, sd, x, y, dtype=torch.float64, **dict(hyper, activation=rec))
    return min(margins)


@pytest.mark.parametrize("n", [32, 33, 48, 64, 128, 256, 512, 1024, 2048])
def test_fno1d_resolutions(n):
    import resolution_pde_b200 as R
    hyper = dict(activation="relu", n_blocks=4)
    for seed in (n, n + 10000, n + 20000):
        torch.manual_seed(seed)
        m = R.FNO1d(1, 1, modes=16, width=32)
        with torch.no_grad():
            for b in m.fno_blocks:
                b.spectral_conv.weights1.mul_(20.0)
        x, y = torch.randn(3, 1, n), torch.randn(3, 1, n)
        if _relu_margin("FNO1d", m, x, y, hyper) > 1e-6:   # seed 512 draws a pre-activation at 1.2e-8 of the rms
            break
    _compare_model("FNO1d", m, x, y, hyper)


def test_fno1d_relu_kink_seed_stays_on_record():
    """The draw test_fno1d_resolutions[512] skips (seed 512) kept visible: the fp64 oracle has a ReLU input at ~1e-8 of the
    rms there, so the mask of that one element is a coin toss in fp32.  The forward pass is continuous at the kink and is held
    to 1e-5; the gradients are compared too, and a miss is reported as an expected failure with the measured numbers."""
    import resolution_pde_b200 as R
    hyper = dict(activation="relu", n_blocks=4)
    torch.manual_seed(512)
    m = R.FNO1d(1, 1, modes=16, width=32)
    with torch.no_grad():
        for b in m.fno_blocks:
            b.spectral_conv.weights1.mul_(20.0)
    x, y = torch.randn(3, 1, 512), torch.randn(3, 1, 512)
    margin = _relu_margin("FNO1d", m, x, y, hyper)
    assert margin < 1e-6, margin      # if this draw ever stops being a kink case it belongs back in the sweep
    sd = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    pred_o, loss_o, g_o = _oracle_step("FNO1d", sd, x, y, **hyper)
    out, loss, gx, g = _step(m.to(DEV), x.to(DEV), y.to(DEV))
    assert rel(out, pred_o) < TOL, rel(out, pred_o)
    worst = max((rel(g[k], v), k) for k, v in g_o.items() if v.abs().max() > 1e-8)
    if worst[0] >= TOL:
        pytest.xfail(f"ReLU mask flip at margin {margin:.2e} of the rms: worst gradient {worst[1]} differs by {worst[0]:.2e}")


@pytest.mark.parametrize("hw", [(32, 32), (64, 64), (64, 128), (128, 128), (33, 47)])
def test_fno2d_resolutions(hw):
    import resolution_pde_b200 as R
    torch.manual_seed(sum(hw))
    m = R.FNO2d(1, 1, modes1=12, modes2=12, width=16)
    with torch.no_grad():
        for b in m.fno_blocks:
            b.spectral_conv.weights1.mul_(10.0)
            b.spectral_conv.weights2.mul_(10.0)
    x, y = torch.randn(2, 1, *hw), torch.randn(2, 1, *hw)
    _compare_model("FNO2d", m, x, y, dict(activation="gelu", n_blocks=4))


@pytest.mark.parametrize("n", [32, 33, 64, 256, 512, 1024])
def test_ffno1d_resolutions(n):
    import resolution_pde_b200 as R
    torch.manual_seed(n + 1)
    m = R.FFNO1D(1, 1, width=32, n_layers=2, n_modes=64, factor=4, ff_weight_norm=True, n_ff_layers=3, layer_norm=True,
                 activation="gelu")
    x, y = torch.randn(4, 1, n), torch.randn(4, 1, n)
    _compare_model("FFNO1D", m, x, y, dict(n_layers=2, n_modes=64, n_ff_layers=3, layer_norm=True, activation="gelu"))


@pytest.mark.parametrize("mn", [(32, 32), (64, 64), (48, 80), (128, 128), (31, 34)])
def test_ffno2d_resolutions(mn):
    import resolution_pde_b200 as R
    torch.manual_seed(sum(mn) + 2)
    m = R.FFNO2D(1, 1, width=32, n_layers=2, n_modes=64, factor=4, ff_weight_norm=True, n_ff_layers=3, layer_norm=True)
    x, y = torch.randn(2, 1, *mn), torch.randn(2, 1, *mn)
    _compare_model("FFNO2D", m, x, y, dict(n_layers=2, n_modes=64, n_ff_layers=3, layer_norm=True, use_grid=True))


def test_ffno2d_no_layernorm_and_lowpass():
    import resolution_pde_b200 as R
    torch.manual_seed(5)
    m = R.FFNO2D(2, 3, width=16, n_layers=2, n_modes=5, factor=2, n_ff_layers=2, layer_norm=False, use_grid=False)
    x, y = torch.randn(2, 2, 24, 20), torch.randn(2, 3, 24, 20)
    _compare_model("FFNO2D", m, x, y, dict(n_layers=2, n_modes=5, n_ff_layers=2, layer_norm=False, use_grid=False))
    m = R.FFNO2D(1, 1, width=16, n_layers=1, n_modes=5, factor=2, n_ff_layers=2, layer_norm=True, mode="low-pass")
    x, y = torch.randn(2, 1, 24, 20), torch.randn(2, 1, 24, 20)
    _compare_model("FFNO2D", m, x, y, dict(n_layers=1, n_modes=5, n_ff_layers=2, layer_norm=True, mode="low-pass"))


# ---------------------------------------------------------------------------------------------------------
# reference error behaviour and quirks
# ---------------------------------------------------------------------------------------------------------
def test_fno_modes_exceed_spectrum_raises():
    import resolution_pde_b200 as R
    with pytest.raises(RuntimeError):
        R.SpectralConv1d(2, 2, 20).to(DEV)(torch.randn(1, 2, 32, device=DEV))
    with pytest.raises(RuntimeError):
        R.SpectralConv2d(2, 2, 4, 10).to(DEV)(torch.randn(1, 2, 16, 16, device=DEV))


def test_ffno_mode_handling():
    import resolution_pde_b200 as R
    x = torch.randn(1, 16, 4, device=DEV)
    with pytest.raises(ValueError):
        R.FSpectralConv1d(4, 3, mode="bogus").to(DEV)(x)
    m = R.FSpectralConv2d(4, 3, mode="bogus").to(DEV)       # silently zero spectrum (reference has no else branch)
    assert m.forward_fourier(torch.randn(1, 8, 8, 4, device=DEV)).abs().max().item() == 0.0
    b, f = R.FSpectralConv2d(4, 3).to(DEV)(torch.randn(1, 8, 8, 4, device=DEV))
    assert f is None and b.shape == (1, 8, 8, 4)
    b, n = R.FSpectralConv1d(4, 3).to(DEV)(x)
    assert n is None and b.shape == (1, 16, 4)


def test_weight_modes_beyond_effective_get_zero_grad():
    import resolution_pde_b200 as R
    m = R.FSpectralConv1d(4, 40, n_ff_layers=2).to(DEV)     # N=16 -> 9 effective modes
    out, _ = m(torch.randn(2, 16, 4, device=DEV))
    out.sum().backward()
    g = m.fourier_weight[0].grad
    assert g[:, :, 9:].abs().max().item() == 0.0 and g[:, :, :9].abs().max().item() > 0.0


# ---------------------------------------------------------------------------------------------------------
# known-answer and property tests (no oracle needed; run at full benchmark sizes)
# ---------------------------------------------------------------------------------------------------------
def test_single_mode_known_answer():
    import resolution_pde_b200 as R
    n, k = 64, 5
    m = R.SpectralConv1d(2, 3, 8).to(DEV)
    t = torch.arange(n, device=DEV, dtype=torch.float64)
    x = torch.zeros(1, 2, n, device=DEV)
    x[0, 1] = torch.cos(2 * np.pi * k * t / n).float()
    y = m(x)
    w = m.weights1.detach()[1, :, k].to(torch.complex128)                       # (out,)
    want = (w[:, None] * torch.exp(2j * np.pi * k * t / n)[None, :]).real       # Re(W e^{i theta})
    assert rel(y[0], want) < 1e-5


def test_identity_weights_nyquist_roundtrip():
    import resolution_pde_b200 as R
    n = 32
    m = R.FSpectralConv1d(4, n // 2 + 1, n_ff_layers=1).to(DEV)
    with torch.no_grad():
        w = torch.zeros_like(m.fourier_weight[0])
        w[range(4), range(4), :, 0] = 1.0
        m.fourier_weight[0].copy_(w)
    x = torch.randn(3, n, 4, device=DEV)
    assert rel(m.forward_fourier(x), x) < 1e-5    # all modes kept + identity mixing = exact identity


def test_lowpass_matches_fft_truncation_large():
    import resolution_pde_b200 as R
    n, k = 2048, 64
    m = R.FSpectralConv1d(8, k, mode="low-pass", n_ff_layers=1).to(DEV)
    x = torch.randn(4, n, 8, device=DEV)
    xf = torch.fft.rfft(x.double().cpu(), dim=1, norm="ortho")
    xf[:, k:] = 0
    want = torch.fft.irfft(xf, n=n, dim=1, norm="ortho")
    assert rel(m.forward_fourier(x), want) < 1e-5


def test_parseval_ortho_full_size():
    """With every mode kept and identity mixing the 2-D factorised pass returns 2x (two axes): checks the C3-sized
    path (B=16, 256x256, C=64 is exercised by the bench; here 4x256x256x64 keeps the test fast)."""
    import resolution_pde_b200 as R
    m = R.FSpectralConv2d(64, 129, n_ff_layers=1).to(DEV)
    with torch.no_grad():
        for p in m.fourier_weight:
            w = torch.zeros_like(p)
            w[range(64), range(64), :, 0] = 1.0
            p.copy_(w)
    x = torch.randn(4, 256, 256, 64, device=DEV)
    assert rel(m.forward_fourier(x), 2 * x) < 1e-5


def test_adjoint_identity_ffno2d_full_size():
    """<A x, y> = <x, A^T y> for the factorised Fourier pass at C3 layer size (backward kernels are the exact
    transposes of the forward ones)."""
    import resolution_pde_b200 as R
    m = R.FSpectralConv2d(64, 64, n_ff_layers=1).to(DEV)
    x = torch.randn(4, 256, 256, 64, device=DEV, requires_grad=True)
    y = torch.randn(4, 256, 256, 64, device=DEV)
    ax = m.forward_fourier(x)
    lhs = (ax.double() * y.double()).sum()
    (aty,) = torch.autograd.grad(ax, x, y)
    rhs = (x.detach().double() * aty.double()).sum()
    assert abs(lhs.item() - rhs.item()) < 1e-5 * max(1.0, abs(lhs.item()))


def test_linearity_fno2d_spectral_full_size():
    import resolution_pde_b200 as R
    m = R.SpectralConv2d(32, 32, 12, 12).to(DEV)
    a, b = torch.randn(4, 32, 256, 256, device=DEV), torch.randn(4, 32, 256, 256, device=DEV)
    with torch.no_grad():
        assert rel(m(a + 2 * b), m(a) + 2 * m(b)) < 1e-5


def test_matmul_engine_vs_fp64():
    import resolution_pde_b200 as R
    torch.manual_seed(0)
    for (M, N, K, ta, tb) in [(130, 70, 33, 0, 0), (64, 64, 64, 1, 0), (257, 129, 1000, 0, 1), (16, 300, 5, 1, 1),
                              (1, 1, 1, 0, 0), (512, 256, 4096, 0, 0)]:
        a = torch.randn((K, M) if ta else (M, K), device=DEV)
        b = torch.randn((N, K) if tb else (K, N), device=DEV)
        got = R.ops.matmul(a, b, bool(ta), bool(tb))
        want = (a.double().T if ta else a.double()) @ (b.double().T if tb else b.double())
        # 3xTF32 chains are cut at 256 k: the accumulator-truncation bias of the tensor core stays near 2e-6 (fp32 FMA: < 1e-6)
        assert rel(got, want) < (4e-6 if R.get_precision() == "3xtf32" else 2e-6), (M, N, K, ta, tb, rel(got, want))


def test_weight_grad_long_reduction_accuracy():
    """dW of a Linear over T = 2^18 rows: split-K keeps fp32 accumulation chains short enough for 1e-5."""
    import resolution_pde_b200 as R
    torch.manual_seed(1)
    lin = R.Linear(64, 128).to(DEV)
    x = torch.randn(1 << 18, 64, device=DEV)
    g = torch.randn(1 << 18, 128, device=DEV)
    lin(x).backward(g)
    want_w = g.double().T @ x.double()
    assert rel(lin.weight.grad, want_w) < TOL
    assert rel(lin.bias.grad, g.double().sum(0)) < TOL


def test_no_grad_inference_and_eval_sweep():
    """Same weights evaluated at every resolution of the super-resolution sweep (resize_utils.py:27-43)."""
    import resolution_pde_b200 as R
    from oracle import reference_port as rp
    torch.manual_seed(3)
    m = R.FFNO1D(1, 1, width=16, n_layers=2, n_modes=16, n_ff_layers=2, layer_norm=True, ff_weight_norm=True).to(DEV)
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    for n in (32, 64, 128, 256, 512, 1024):
        x = torch.randn(2, 1, n)
        with torch.no_grad():
            got = m(x.to(DEV))
        want = rp.ffno1d_forward({k: v.double() for k, v in sd.items()}, x.double(), n_layers=2, n_modes=16,
                                 n_ff_layers=2, layer_norm=True)
        assert rel(got, want) < TOL, n


# ---------------------------------------------------------------------------------------------------------
# graph-captured step == eager step (same kernels, same order), one graph per resolution
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind", ["FNO1d", "FFNO2D"])
def test_graphed_step_matches_eager(kind):
    import resolution_pde_b200 as R
    torch.manual_seed(7)
    if kind == "FNO1d":
        m = R.FNO1d(1, 1, modes=8, width=16).to(DEV)
        shapes = [(3, 1, 64), (2, 1, 128)]
    else:
        m = R.FFNO2D(1, 1, width=16, n_layers=2, n_modes=8, n_ff_layers=2, layer_norm=True, ff_weight_norm=True).to(DEV)
        shapes = [(2, 1, 32, 32), (1, 1, 48, 40)]
    loss_fn = R.RelativeL2Loss()
    gstep = R.GraphedTrainStep(m, loss_fn)
    for rep in range(2):                       # second round replays the cached graphs with new data
        for shp in shapes:
            x, y = torch.randn(*shp, device=DEV), torch.randn(*shp, device=DEV)
            lg = gstep(x, y).clone()
            gg = {k: p.grad.clone() for k, p in m.named_parameters()}
            m.zero_grad(set_to_none=True)
            le = loss_fn(m(x), y)
            le.backward()
            torch.cuda.synchronize()
            assert abs(lg.item() - le.item()) <= 1e-6 * max(1.0, abs(le.item()))
            for k, p in m.named_parameters():
                # same kernels in the same order; parameter gradients that leave their CTAs as fp32 atomics (bias / weight sums
                # of the fused MLP and the 1x1 convs) differ from run to run by the order of those adds: a few 1e-6
                assert rel(gg[k], p.grad) < 5e-6 or p.grad.abs().max() < 1e-7, (k, rel(gg[k], p.grad))
            m.zero_grad(set_to_none=True)
