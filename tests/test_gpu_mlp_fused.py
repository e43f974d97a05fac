"""GPU (-m gpu): the fused projection MLP (rpde_mlp_fwd / rpde_mlp_bwd, csrc/fno2d_mma.cuh; fno_blocks.py:35-45, 73-83)
against a torch fp64 restatement of the same three lines (conv1x1 -> gelu -> conv1x1), forward and every gradient, through
the autograd wrapper (ops.mlp_gelu) and directly through the C ABI.  Shapes cover both register tilings (I <= 32, I <= 64),
hidden widths 32 .. 256 incl. ones that are padded, 1 .. 4 outputs, channel counts that are not multiples of 8 and tile
counts that do not fill the last iteration.  Tolerance 1e-5 relative L2 (helpers.TOL)."""
import pytest
import torch
import torch.nn.functional as F

from helpers import TOL, rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

SHAPES = [  # B, I, Hd, O, spatial
    (16, 32, 128, 1, (64, 64)),     # C4 projection at 64^2
    (2, 64, 256, 1, (1024,)),       # C2 projection
    (3, 20, 100, 3, (96,)),         # padded channels and hidden rows, 9 tiles against 2 per iteration
    (1, 8, 32, 2, (32,)),           # one tile, eight point tiles per iteration
    (2, 64, 64, 4, (160,)),
    (1, 33, 128, 1, (8, 8)),
    (5, 1, 128, 1, (64,)),
]


def _make(B, I, Hd, O, spatial, seed):
    g = torch.Generator().manual_seed(seed)
    nd = len(spatial)
    x = torch.randn((B, I) + spatial, generator=g)
    w1 = torch.randn((Hd, I) + (1,) * nd, generator=g) / I ** 0.5
    b1 = torch.randn(Hd, generator=g)
    w2 = torch.randn((O, Hd) + (1,) * nd, generator=g) / Hd ** 0.5
    b2 = torch.randn(O, generator=g)
    gy = torch.randn((B, O) + spatial, generator=g)
    return x, w1, b1, w2, b2, gy


def _oracle(x, w1, b1, w2, b2):
    conv = F.conv1d if x.dim() == 3 else F.conv2d
    return conv(F.gelu(conv(x, w1, b1)), w2, b2)


@pytest.mark.parametrize("B,I,Hd,O,spatial", SHAPES)
@pytest.mark.parametrize("bias", [True, False])
def test_mlp_fused_forward_and_gradients(B, I, Hd, O, spatial, bias):
    import resolution_pde_b200 as R
    from resolution_pde_b200 import ops
    R.set_precision("fp32")
    S = 1
    for v in spatial:
        S *= v
    assert R._lib.load().rpde_mlp_supported(I, Hd, O, S) == 1
    src = list(_make(B, I, Hd, O, spatial, B * 100 + I + Hd))
    gy = src.pop()
    if not bias:
        src[2] = src[4] = None
    lo = [None if t is None else t.clone().double().requires_grad_(True) for t in src]
    yo = _oracle(*lo)
    yo.backward(gy.double())
    for fused in (True, False):
        old, ops._mlp_fused = ops._mlp_fused, fused
        try:
            leaves = [None if t is None else t.clone().to(DEV).requires_grad_(True) for t in src]
            n0 = R._lib.load().rpde_launch_count()
            y = ops.mlp_gelu(*leaves)
            launches = R._lib.load().rpde_launch_count() - n0
            y.backward(gy.to(DEV))
        finally:
            ops._mlp_fused = old
        assert (launches == 1) if fused else (launches >= 2)
        assert rel(y, yo) < TOL, (fused, rel(y, yo))
        for name, a, o in zip(("x", "w1", "b1", "w2", "b2"), leaves, lo):
            if o is not None:
                assert rel(a.grad, o.grad) < TOL, (fused, name, rel(a.grad, o.grad))


def test_mlp_fused_without_input_gradient_and_under_no_grad():
    from resolution_pde_b200 import ops
    x, w1, b1, w2, b2, gy = [t.to(DEV) for t in _make(2, 32, 128, 1, (32, 32), 3)]
    with torch.no_grad():
        y0 = ops.mlp_gelu(x, w1, b1, w2, b2)
    w1g = w1.clone().requires_grad_(True)
    y1 = ops.mlp_gelu(x, w1g, b1, w2, b2)          # x needs no gradient: dx is not computed
    assert torch.equal(y0, y1.detach())
    y1.backward(gy)
    lo = [t.double().cpu() for t in (x, w1, b1, w2, b2)]
    lo[1].requires_grad_(True)
    _oracle(*lo).backward(gy.double().cpu())
    assert rel(w1g.grad, lo[1].grad) < TOL


def test_mlp_unsupported_shapes_report_and_fall_back():
    import resolution_pde_b200 as R
    from resolution_pde_b200 import ops
    lib = R._lib.load()
    assert lib.rpde_mlp_supported(32, 128, 1, 33) == 0      # S % 32 != 0
    assert lib.rpde_mlp_supported(96, 128, 1, 64) == 0      # more than 64 input channels
    assert lib.rpde_mlp_supported(32, 96, 1, 64) == 0       # hidden width rounds to 96
    assert lib.rpde_mlp_supported(32, 128, 5, 64) == 0
    x, w1, b1, w2, b2, gy = _make(2, 16, 64, 1, (33,), 9)
    lo = [t.clone().double().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    _oracle(*lo).backward(gy.double())
    leaves = [t.clone().to(DEV).requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    ops.mlp_gelu(*leaves).backward(gy.to(DEV))
    for a, o in zip(leaves, lo):
        assert rel(a.grad, o.grad) < TOL
