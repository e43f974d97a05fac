"""GPU tests of individual C-ABI entry points (include/rpde_b200.h) against plain torch fp64 references of the same op:
fused FeedForward epilogues (stored gelu', fused bias gradients), few-channel streaming Linear kernels, LayerNorm
backward column sums, DFT table layouts.  Tolerances: 1e-5 relative L2 (fp32 / 3xTF32 engines)."""
import ctypes
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-5


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


@pytest.fixture(params=["fp32", "3xtf32"])
def lib(request):
    import resolution_pde_b200 as R
    R.set_precision(request.param)
    R.ops.matmul(torch.randn(8, 8, device=DEV), torch.randn(8, 8, device=DEV))  # pushes the engine choice to the C side
    yield R._lib.load()
    R.set_precision("fp32")
    R.ops.matmul(torch.randn(8, 8, device=DEV), torch.randn(8, 8, device=DEV))


def _st():
    return torch.cuda.current_stream().cuda_stream


def _gelu_grad(z):
    return 0.5 * (1 + torch.erf(z / math.sqrt(2))) + z * torch.exp(-0.5 * z * z) / math.sqrt(2 * math.pi)


@pytest.mark.parametrize("T,i,o", [(4096, 64, 256), (1000, 32, 96), (8192, 256, 64)])
def test_linear_fwd_saves_gelu_grad(lib, T, i, o):
    torch.manual_seed(T + i)
    x = torch.randn(T, i, device=DEV)
    w = torch.randn(o, i, device=DEV) / math.sqrt(i)
    b = torch.randn(o, device=DEV)
    y = torch.empty(T, o, device=DEV)
    g = torch.empty(T, o, device=DEV)
    rc = lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), g.data_ptr(), T, i, o, 0x101, 0, _st())
    assert rc == 0, lib.rpde_last_error()
    z = x.double() @ w.double().T + b.double()
    assert rel(y, torch.nn.functional.gelu(z)) < TOL
    assert rel(g, _gelu_grad(z)) < TOL


@pytest.mark.parametrize("T,i,o", [(4096, 256, 64), (3000, 96, 32), (8192, 256, 256)])
def test_linear_bwd_input_saved_factor_and_bias_grad(lib, T, i, o):
    torch.manual_seed(T + o)
    dy = torch.randn(T, o, device=DEV)
    w = torch.randn(o, i, device=DEV) / math.sqrt(o)
    f = torch.rand(T, i, device=DEV) + 0.1            # stored act'(z) of the producing layer
    dx = torch.empty(T, i, device=DEV)
    db = torch.full((i,), 7.0, device=DEV)            # must be overwritten
    rc = lib.rpde_linear_bwd_input(dy.data_ptr(), w.data_ptr(), f.data_ptr(), 3, dx.data_ptr(), 0, db.data_ptr(), T, i, o, 0, _st())
    assert rc == 0, lib.rpde_last_error()
    ref = (dy.double() @ w.double()) * f.double()
    assert rel(dx, ref) < TOL
    assert rel(db, ref.sum(0)) < TOL


@pytest.mark.parametrize("cin", [1, 2, 3, 4])
def test_few_channel_linear_streaming_kernels(lib, cin):
    T, o = 5000, 64
    torch.manual_seed(cin)
    x = torch.randn(T, cin, device=DEV)
    w = torch.randn(o, cin, device=DEV)
    b = torch.randn(o, device=DEV)
    y = torch.empty(T, o, device=DEV)
    assert lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), None, T, cin, o, 0, 0, _st()) == 0
    assert rel(y, x.double() @ w.double().T + b.double()) < TOL
    dy = torch.randn(T, o, device=DEV)
    dw = torch.empty(o, cin, device=DEV)
    dbv = torch.empty(o, device=DEV)
    assert lib.rpde_linear_bwd_params(dy.data_ptr(), x.data_ptr(), dw.data_ptr(), dbv.data_ptr(), T, cin, o, 0, _st()) == 0
    assert rel(dw, dy.double().T @ x.double()) < TOL
    assert rel(dbv, dy.double().sum(0)) < TOL


def test_layernorm_bwd_column_sums(lib):
    T, C = 3000, 64
    torch.manual_seed(5)
    x = torch.randn(T, C, device=DEV, dtype=torch.float64, requires_grad=True)
    gamma = torch.randn(C, device=DEV, dtype=torch.float64, requires_grad=True)
    beta = torch.randn(C, device=DEV, dtype=torch.float64, requires_grad=True)
    dy = torch.randn(T, C, device=DEV, dtype=torch.float64)
    torch.nn.functional.layer_norm(x, (C,), gamma, beta, 1e-5).backward(dy)
    xf, gf, dyf = x.detach().float(), gamma.detach().float(), dy.float()
    mean = xf.mean(1).contiguous()
    rstd = (1.0 / torch.sqrt(xf.double().var(1, unbiased=False) + 1e-5)).float().contiguous()
    dx = torch.empty(T, C, device=DEV)
    dg, dbt, dxs = (torch.empty(C, device=DEV) for _ in range(3))
    rc = lib.rpde_layernorm_bwd(dyf.data_ptr(), xf.data_ptr(), gf.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dx.data_ptr(),
                                dg.data_ptr(), dbt.data_ptr(), dxs.data_ptr(), T, C, _st())
    assert rc == 0, lib.rpde_last_error()
    assert rel(dx, x.grad) < TOL and rel(dg, gamma.grad) < TOL and rel(dbt, beta.grad) < TOL
    assert rel(dxs, x.grad.sum(0)) < 1e-4            # a sum of T cancelling terms: absolute accuracy, looser relative bound


@pytest.mark.parametrize("n,k", [(64, 12), (33, 7), (256, 129)])
def test_dft_table_planes_match_interleaved(n, k):
    import resolution_pde_b200 as R
    t1 = R.ops.dft_table(1, n, k, 0, 1.0 / n, DEV).reshape(n, k, 2)
    t4 = R.ops.dft_table(4, n, k, 0, 1.0 / n, DEV).reshape(2, k, n)
    assert torch.equal(t4, t1.permute(2, 1, 0).contiguous())
    # synthesis table reproduces irfft of a truncated spectrum
    torch.manual_seed(n)
    spec = torch.randn(k, dtype=torch.complex128)
    full = torch.zeros(n // 2 + 1, dtype=torch.complex128)
    full[:k] = spec
    want = torch.fft.irfft(full, n=n)
    got = t4[0].double().cpu().T @ spec.real + t4[1].double().cpu().T @ spec.imag
    assert rel(got, want) < 1e-6


@pytest.mark.parametrize("T,C", [(3001, 16), (777, 32), (4099, 64), (1025, 128), (513, 48), (300, 20)])
@pytest.mark.parametrize("with_res", [False, True])
def test_layernorm_fwd_bwd_all_widths(lib, T, C, with_res):
    """C in {16, 32, 64, 128} runs the vectorised kernels (float4 per lane, ragged row tails), other widths the scalar ones."""
    torch.manual_seed(T + C)
    x = torch.randn(T, C, device=DEV, dtype=torch.float64, requires_grad=True) * 3 + 0.5
    x.retain_grad()
    gamma = torch.randn(C, device=DEV, dtype=torch.float64, requires_grad=True)
    beta = torch.randn(C, device=DEV, dtype=torch.float64, requires_grad=True)
    res = torch.randn(T, C, device=DEV, dtype=torch.float64)
    dy = torch.randn(T, C, device=DEV, dtype=torch.float64)
    ref = torch.nn.functional.layer_norm(x, (C,), gamma, beta, 1e-5) + (res if with_res else 0)
    ref.backward(dy)
    xf, gf, bf, rf, dyf = x.detach().float(), gamma.detach().float(), beta.detach().float(), res.float(), dy.float()
    y = torch.empty(T, C, device=DEV)
    mean, rstd = torch.empty(T, device=DEV), torch.empty(T, device=DEV)
    rc = lib.rpde_layernorm_fwd(xf.data_ptr(), rf.data_ptr() if with_res else None, gf.data_ptr(), bf.data_ptr(), y.data_ptr(),
                                mean.data_ptr(), rstd.data_ptr(), T, C, ctypes.c_float(1e-5), _st())
    assert rc == 0, lib.rpde_last_error()
    assert rel(y, ref) < TOL
    assert rel(mean, xf.double().mean(1)) < TOL
    assert rel(rstd, 1.0 / torch.sqrt(xf.double().var(1, unbiased=False) + 1e-5)) < TOL
    dx = torch.empty(T, C, device=DEV)
    dg, dbt, dxs = (torch.empty(C, device=DEV) for _ in range(3))
    rc = lib.rpde_layernorm_bwd(dyf.data_ptr(), xf.data_ptr(), gf.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dx.data_ptr(),
                                dg.data_ptr(), dbt.data_ptr(), dxs.data_ptr(), T, C, _st())
    assert rc == 0, lib.rpde_last_error()
    assert rel(dx, x.grad) < TOL and rel(dg, gamma.grad) < TOL and rel(dbt, beta.grad) < TOL
    assert (dxs.double() - x.grad.sum(0)).abs().max().item() < 1e-4 * x.grad.abs().sum(0).max().item()


@pytest.mark.parametrize("cin,S", [(1, 1024), (2, 512), (3, 64 * 64), (4, 260), (3, 33 * 47), (5, 256)])
def test_lifting_conv_and_bias_grad_reduction(lib, cin, S):
    """Channels-first 1x1 conv with few input channels (streaming kernel when S % 4 == 0 and cin <= 4, contraction engine
    otherwise) and its parameter gradients (the bias gradient is the vectorised channel reduction)."""
    B, O = 3, 32
    torch.manual_seed(cin * 1000 + S)
    x = torch.randn(B, cin, S, device=DEV, dtype=torch.float64)
    w = torch.randn(O, cin, device=DEV, dtype=torch.float64, requires_grad=True)
    b = torch.randn(O, device=DEV, dtype=torch.float64, requires_grad=True)
    dy = torch.randn(B, O, S, device=DEV, dtype=torch.float64)
    ref = torch.einsum("oi,bis->bos", w, x) + b[None, :, None]
    ref.backward(dy)
    xf, wf, bf, dyf = x.float(), w.detach().float().contiguous(), b.detach().float(), dy.float()
    y = torch.empty(B, O, S, device=DEV)
    rc = lib.rpde_conv1x1_fwd(xf.data_ptr(), wf.data_ptr(), bf.data_ptr(), y.data_ptr(), None, B, cin, O, S, 0, 0, _st())
    assert rc == 0, lib.rpde_last_error()
    assert rel(y, ref) < TOL
    dw, db = torch.empty(O, cin, device=DEV), torch.empty(O, device=DEV)
    rc = lib.rpde_conv1x1_bwd_params(dyf.data_ptr(), xf.data_ptr(), dw.data_ptr(), db.data_ptr(), B, cin, O, S, _st())
    assert rc == 0, lib.rpde_last_error()
    assert rel(dw, w.grad) < TOL and rel(db, b.grad) < TOL


@pytest.mark.parametrize("B,I,O,S,bias", [(16, 32, 32, 64 * 64, True), (3, 32, 32, 96, False), (2, 64, 64, 1024, True),
                                          (2, 20, 12, 256, True), (1, 64, 32, 32, True), (5, 1, 128, 64, True),
                                          (2, 128, 1, 64, True), (4, 40, 64, 2048, False)])
def test_conv1x1_param_gradients_streaming_tensor_path(lib, B, I, O, S, bias):
    """rpde_conv1x1_bwd_params with S % 32 == 0 and at most 64 channels per side: one streaming pass on the warp-level tensor
    path (conv1x1_wgrad_mma_kernel, 32 x 32 blocks of (o, i) spread over the warps); wider layers keep the contraction engine."""
    torch.manual_seed(B * 100 + I + O + S)
    x = torch.randn(B, I, S, device=DEV, dtype=torch.float64) + 0.3
    dy = torch.randn(B, O, S, device=DEV, dtype=torch.float64) - 0.2
    want_w = torch.einsum("bos,bis->oi", dy, x)
    want_b = dy.sum((0, 2))
    xf, dyf = x.float(), dy.float()
    for _ in range(2):  # twice: the outputs are accumulated with atomics and must be re-zeroed by the entry point
        dw, db = torch.full((O, I), 7.0, device=DEV), torch.full((O,), 7.0, device=DEV)
        rc = lib.rpde_conv1x1_bwd_params(dyf.data_ptr(), xf.data_ptr(), dw.data_ptr(), db.data_ptr() if bias else None, B, I, O,
                                         S, _st())
        assert rc == 0, lib.rpde_last_error()
        assert rel(dw, want_w) < TOL, rel(dw, want_w)
        if bias:
            assert rel(db, want_b) < TOL, rel(db, want_b)
