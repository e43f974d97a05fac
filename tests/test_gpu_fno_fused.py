"""GPU (-m gpu): the fused FNOBlock1d forward (rpde_fno_block1d_fwd: analysis + mixing, synthesis + bypass + bias + act)
against the fp64 oracle and against the composed path (rpde_conv1x1_fwd + rpde_spectral1d_fwd), forward and gradients
(the backward pass consumes the spectrum the fused kernel stored)."""
import pytest
import torch
import torch.nn.functional as F

from helpers import TOL, rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ENGINES = ["fp32", "3xtf32"]


@pytest.fixture(params=ENGINES)
def engine(request):
    import resolution_pde_b200 as R
    R.set_precision(request.param)
    yield request.param
    R.set_precision("fp32")


@pytest.fixture(autouse=True, params=[False, True], ids=["bwd-composed", "bwd-fused"])
def _fused_backward(request):
    """both backward prologues: the composed one (default) and rpde_fno_block1d_bwd (RPDE_FNO_FUSED_BWD=1)"""
    from resolution_pde_b200 import ops
    old, ops._fno_fused_bwd = ops._fno_fused_bwd, request.param
    yield
    ops._fno_fused_bwd = old


@pytest.fixture
def composed():
    """context: run the composed (unfused) 1-D path"""
    from resolution_pde_b200 import ops

    class _C:
        def __enter__(self):
            self.old, ops._fno_fused = ops._fno_fused, False

        def __exit__(self, *a):
            ops._fno_fused = self.old
    return _C()


def _oracle(x, w, wb, bb, act):
    from oracle import reference_port as rp
    s = rp.spectral_conv1d(x.double(), w.to(torch.complex128))
    if wb is not None:
        s = s + F.conv1d(x.double(), wb.double(), None if bb is None else bb.double())
    return {"none": lambda v: v, "relu": F.relu, "gelu": F.gelu}[act](s)


SHAPES = [  # B, I, O, N, K, bypass, bias, act
    (16, 64, 64, 1024, 16, True, True, "relu"),
    (2, 16, 16, 64, 8, True, True, "gelu"),
    (3, 3, 5, 33, 7, False, False, "none"),
    (2, 8, 96, 200, 5, True, False, "gelu"),
    (1, 70, 33, 47, 24, True, True, "relu"),
    (2, 4, 4, 16, 9, True, True, "none"),          # Nyquist bin kept
    (2, 32, 32, 2048, 1, True, True, "gelu"),
    # >= 4736 rows (b, i): the analysis runs on the warp-level tensor path (csrc/fno2d_mma.cuh), table as one fp32 plane
    # split on the fly at N = 1024, as resident TF32 hi / lo planes at N = 256
    (80, 64, 64, 1024, 16, True, True, "gelu"),
    (160, 32, 32, 256, 12, True, True, "relu"),
]


@pytest.mark.parametrize("B,I,O,N,K,bypass,bias,act", SHAPES)
def test_fused_block_forward_and_gradients(engine, composed, B, I, O, N, K, bypass, bias, act):
    from resolution_pde_b200 import ops
    g = torch.Generator().manual_seed(B * 1000 + N + K)
    x = torch.randn(B, I, N, generator=g)
    w = torch.complex(torch.rand(I, O, K, generator=g), torch.rand(I, O, K, generator=g)) / (I * O) ** 0.5
    wb = torch.randn(O, I, 1, generator=g) / I ** 0.5 if bypass else None
    bb = torch.randn(O, generator=g) if bias else None
    aid = {"none": ops.ACT_NONE, "relu": ops.ACT_RELU, "gelu": ops.ACT_GELU}[act]

    def run(fused):
        leaves = [t.clone().to(DEV).requires_grad_(True) if t is not None else None for t in (x, w, wb, bb)]
        if fused:
            y = ops.fno_spectral(leaves[0], leaves[1], None, leaves[2], leaves[3], aid)
        else:
            with composed:
                y = ops.fno_spectral(leaves[0], leaves[1], None, leaves[2], leaves[3], aid)
        gy = torch.randn(y.shape, generator=torch.Generator().manual_seed(1)).to(DEV)
        y.backward(gy)
        return y.detach(), [None if t is None else t.grad for t in leaves], gy

    yf, gf, gy = run(True)
    yc, gc, _ = run(False)
    lo = [t.clone().double().requires_grad_(True) if t is not None and not t.is_complex() else
          (t.clone().to(torch.complex128).requires_grad_(True) if t is not None else None) for t in (x, w, wb, bb)]
    yo = _oracle(lo[0], lo[1], lo[2], lo[3], act)
    yo.backward(gy.double().cpu())
    assert rel(yf, yo) < TOL and rel(yc, yo) < TOL
    for a, c, o in zip(gf, gc, lo):
        if o is None:
            continue
        assert rel(a, o.grad) < TOL, rel(a, o.grad)
        assert rel(c, o.grad) < TOL


def test_fused_inference_skips_spectrum_and_preactivation():
    from resolution_pde_b200 import ops
    x = torch.randn(4, 32, 256, device=DEV)
    w = torch.complex(torch.rand(32, 32, 12), torch.rand(32, 32, 12)).to(DEV) / 32
    wb, bb = torch.randn(32, 32, 1, device=DEV) / 6, torch.randn(32, device=DEV)
    with torch.no_grad():
        y0 = ops.fno_spectral(x, w, None, wb, bb, ops.ACT_GELU)
    xr = x.clone().requires_grad_(True)
    y1 = ops.fno_spectral(xr, w, None, wb, bb, ops.ACT_GELU)
    # same kernels either way; with few rows the analysis splits each row over several CTAs whose partial sums meet through
    # fp32 atomics, so two runs agree to the order of those adds, not bit for bit
    assert rel(y0, y1.detach()) < 1e-6


def test_composed_path_still_matches_fixture(composed):
    """the unfused 1-D path stays covered: the FNO1d golden fixture through it"""
    from helpers import build_candidate, load_fixture
    z, params, grads, hyper = load_fixture("fno1d_n64.npz")
    m, _ = build_candidate("fno1d_n64.npz", params, hyper)
    m.load_state_dict(params)
    m = m.to(DEV)
    with composed:
        out = m(torch.from_numpy(z["x"]).to(DEV))
    assert rel(out, z["out"]) < TOL


@pytest.mark.parametrize("force", [True, False])
@pytest.mark.parametrize("B,I,O,N,K", [(3, 24, 40, 256, 9), (2, 64, 64, 1024, 16)])
def test_fused_input_gradient_switch(force, B, I, O, N, K):
    """dx from one kernel (tensor-path synthesis kernel in adjoint mode: analysis adjoint + wb^T dpre) or from the analysis-adjoint
    contraction plus rpde_conv1x1_bwd_input: the automatic choice (ops._use_fused_dx) may take either, both must match the oracle."""
    from resolution_pde_b200 import ops
    g = torch.Generator().manual_seed(N + K)
    x = torch.randn(B, I, N, generator=g)
    w = torch.complex(torch.rand(I, O, K, generator=g), torch.rand(I, O, K, generator=g)) / (I * O) ** 0.5
    wb = torch.randn(O, I, 1, generator=g) / I ** 0.5
    bb = torch.randn(O, generator=g)
    gy = torch.randn(B, O, N, generator=g)
    old, ops._fno_fused_dx = ops._fno_fused_dx, force
    try:
        leaves = [t.clone().to(DEV).requires_grad_(True) for t in (x, w, wb, bb)]
        y = ops.fno_spectral(leaves[0], leaves[1], None, leaves[2], leaves[3], ops.ACT_GELU)
        y.backward(gy.to(DEV))
    finally:
        ops._fno_fused_dx = old
    lo = [x.double().requires_grad_(True), w.to(torch.complex128).requires_grad_(True), wb.double().requires_grad_(True),
          bb.double().requires_grad_(True)]
    yo = _oracle(lo[0], lo[1], lo[2], lo[3], "gelu")
    yo.backward(gy.double())
    assert rel(y, yo) < TOL
    for a, o in zip(leaves, lo):
        assert rel(a.grad, o.grad) < TOL, rel(a.grad, o.grad)
