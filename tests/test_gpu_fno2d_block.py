"""GPU (-m gpu): the fused FNOBlock2d (rpde_fno_block2d_fwd / _bwd) at the block level against the fp64 oracle
(spectral_conv2d + 1x1 bypass conv + bias + activation, fno_blocks.py:63-71), forward and every gradient.

The image-row widths 64 / 128 / 256 with at most 32 channels take the tensor-path W-axis kernels (csrc/fno2d_mma.cuh,
mma.sync 3xTF32); the other shapes take the fp32-FMA kernels.  Cases cover channel counts that are not multiples of the
MMA tile (padding), 2*m2 not a multiple of 8, a row count that does not fill the last iteration, no bypass / no bias.
Tolerance: 1e-5 relative L2 (helpers.TOL), the fp32 bar of north_star."""
import pytest
import torch
import torch.nn.functional as F

from helpers import TOL, rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _oracle(x, w1, w2, wb, bb, act):
    from oracle import reference_port as rp
    s = rp.spectral_conv2d(x, w1, w2)
    if wb is not None:
        s = s + F.conv2d(x, wb, bb)
    return {"none": lambda v: v, "relu": F.relu, "gelu": F.gelu}[act](s)


SHAPES = [  # B, I, O, H, W, m1, m2, bypass, bias, act
    (2, 32, 32, 64, 64, 12, 12, True, True, "gelu"),      # C4 block at 64^2
    (1, 32, 32, 24, 256, 5, 12, True, True, "gelu"),      # one row per iteration
    (2, 32, 32, 9, 128, 4, 12, True, True, "relu"),       # 18 rows, two per iteration
    (1, 16, 16, 3, 64, 1, 5, True, True, "gelu"),         # 3 rows against 4 per iteration, 2*m2 = 10 -> padded to 16
    (2, 20, 12, 10, 128, 3, 7, True, False, "none"),      # channels not multiples of 8
    (1, 3, 5, 6, 64, 2, 16, False, False, "none"),        # no bypass, 2*m2 = 32
    (2, 32, 32, 16, 256, 6, 16, True, True, "gelu"),
    (1, 32, 32, 5, 512, 2, 12, True, True, "gelu"),       # long rows: two 256-point chunks per image row, whole table resident
    (1, 16, 16, 3, 1024, 1, 5, True, True, "relu"),       # four chunks; W-axis analysis with the fp32 table plane
    (1, 8, 8, 12, 96, 3, 6, True, True, "gelu"),          # W = 96: the fp32-FMA kernels
    (1, 40, 40, 8, 64, 3, 6, True, True, "relu"),         # 40 channels: the fp32-FMA kernels
]


@pytest.mark.parametrize("B,I,O,H,W,m1,m2,bypass,bias,act", SHAPES)
def test_fno_block2d_forward_and_gradients(B, I, O, H, W, m1, m2, bypass, bias, act):
    import resolution_pde_b200 as R
    from resolution_pde_b200 import ops
    R.set_precision("fp32")
    lib = R._lib.load()
    assert lib.rpde_fno_block2d_supported(B, I, O, H, W, m1, m2) == 1
    g = torch.Generator().manual_seed(B * 1000 + H * 7 + W + m2)
    x = torch.randn(B, I, H, W, generator=g)
    sc = 1.0 / (I * O) ** 0.5
    w1 = torch.complex(torch.rand(I, O, m1, m2, generator=g), torch.rand(I, O, m1, m2, generator=g)) * sc
    w2 = torch.complex(torch.rand(I, O, m1, m2, generator=g), torch.rand(I, O, m1, m2, generator=g)) * sc
    wb = torch.randn(O, I, 1, 1, generator=g) / I ** 0.5 if bypass else None
    bb = torch.randn(O, generator=g) if bias else None
    aid = {"none": ops.ACT_NONE, "relu": ops.ACT_RELU, "gelu": ops.ACT_GELU}[act]
    src = (x, w1, w2, wb, bb)

    leaves = [None if t is None else t.clone().to(DEV).requires_grad_(True) for t in src]
    y = ops.fno_spectral(leaves[0], leaves[1], leaves[2], leaves[3], leaves[4], aid)
    gy = torch.randn(y.shape, generator=torch.Generator().manual_seed(1))
    y.backward(gy.to(DEV))

    lo = [None if t is None else (t.clone().to(torch.complex128) if t.is_complex() else t.clone().double()).requires_grad_(True)
          for t in src]
    yo = _oracle(*lo, act)
    yo.backward(gy.double())
    assert rel(y, yo) < TOL, rel(y, yo)
    for name, a, o in zip(("x", "w1", "w2", "wb", "bb"), leaves, lo):
        if o is None:
            continue
        assert rel(a.grad, o.grad) < TOL, (name, rel(a.grad, o.grad))


def test_fno_block2d_inference_matches_training_forward():
    """no_grad forward (no spectrum, no pre-activation stored) gives the same output as the training forward"""
    from resolution_pde_b200 import ops
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 32, 32, 128, generator=g).to(DEV)
    w1 = (torch.complex(torch.rand(32, 32, 6, 12, generator=g), torch.rand(32, 32, 6, 12, generator=g)) / 32).to(DEV)
    w2 = (torch.complex(torch.rand(32, 32, 6, 12, generator=g), torch.rand(32, 32, 6, 12, generator=g)) / 32).to(DEV)
    wb = (torch.randn(32, 32, 1, 1, generator=g) / 32 ** 0.5).to(DEV)
    bb = torch.randn(32, generator=g).to(DEV)
    with torch.no_grad():
        y0 = ops.fno_spectral(x, w1, w2, wb, bb, ops.ACT_GELU)
    y1 = ops.fno_spectral(x.clone().requires_grad_(True), w1, w2, wb, bb, ops.ACT_GELU)
    assert rel(y0, y1.detach()) < 1e-6      # bit-equal unless the W-axis analysis split its rows over CTAs (atomics)
