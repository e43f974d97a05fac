"""GPU tests of the tcgen05/TMA contraction engine (gemm_tc.cuh) through the C ABI test hook rpde_matmul:
every operand-major combination, ragged sizes, split-free and long-K shapes, both precision variants,
against an fp64 product of the same inputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu

# 3xTF32: the tensor core truncates its fp32 accumulator after every UMMA (2.2e-8 relative per UMMA, measured); chains are
# cut at 256 k (96 UMMAs) and chunk sums are added in registers, which bounds the bias near 2e-6 for any K.
TOL = {"3xtf32": 4e-6, "tf32": 2e-3, "fp32": 2e-6}
SHAPES = [(256, 256, 256), (128, 64, 32), (200, 136, 100), (1000, 72, 260), (4096, 256, 64), (4096, 64, 256), (384, 192, 4096),
          (130, 260, 20000)]


@pytest.fixture(autouse=True)
def _restore():
    import resolution_pde_b200 as R
    yield
    R.set_precision("fp32")


@pytest.mark.parametrize("prec", ["3xtf32", "tf32"])
@pytest.mark.parametrize("ta,tb", [(0, 1), (0, 0), (1, 1), (1, 0)])
@pytest.mark.parametrize("shape", SHAPES)
def test_matmul_majors(shape, ta, tb, prec):
    import resolution_pde_b200 as R
    M, N, K = shape
    torch.manual_seed(M + N + K)
    a = torch.randn((K, M) if ta else (M, K), device="cuda")
    b = torch.randn((N, K) if tb else (K, N), device="cuda")
    R.set_precision(prec)
    lib = R._lib.load()
    c0 = lib.rpde_tc_launch_count()
    out = R.ops.matmul(a, b, bool(ta), bool(tb))
    used = lib.rpde_tc_launch_count() - c0
    ad, bd = a.double(), b.double()
    ref = (ad.t() if ta else ad) @ (bd.t() if tb else bd)
    err = ((out.double() - ref).norm() / ref.norm()).item()
    assert err < TOL[prec], (err, used)
    # strides that are multiples of 16 bytes must take the tensor-core engine
    lead_a, lead_b = (M if ta else K), (K if tb else N)
    if lead_a % 4 == 0 and lead_b % 4 == 0:
        assert used == 1


def test_unaligned_falls_back_to_cuda_cores():
    import resolution_pde_b200 as R
    R.set_precision("3xtf32")
    lib = R._lib.load()
    a = torch.randn(130, 67, device="cuda")
    b = torch.randn(67, 129, device="cuda")
    c0 = lib.rpde_tc_launch_count()
    out = R.ops.matmul(a, b)
    assert lib.rpde_tc_launch_count() == c0
    ref = a.double() @ b.double()
    assert ((out.double() - ref).norm() / ref.norm()).item() < 2e-6
