"""GPU (-m gpu): the non-contraction data-preparation kernels (csrc/dataprep.cu) against the oracle's transcription of the
reference's host code: global mean / std normaliser statistics over mixed-resolution samples, naive stride
down-sampling (bit-exact: it is a gather), and the radial-band sums of the frequency-error decomposition."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_global_mean_std_matches_reference_procedure():
    from oracle import spectral_ops_port as op
    import resolution_pde_b200.spectral_ops as so
    g = torch.Generator().manual_seed(5)
    samples = [3.0 + 2.0 * torch.randn(1, n, n, generator=g) for n in (16, 16, 24, 32, 33)] + [torch.randn(7, generator=g)]
    m_ref, s_ref = op.global_mean_std(samples)
    m, s = so.global_mean_std([t.to(DEV) for t in samples])
    assert abs(m - m_ref) < 1e-6 * max(1.0, abs(m_ref)) and abs(s - s_ref) < 1e-6 * s_ref
    big = 0.5 + torch.randn(4, 1, 256, 256, generator=g)
    m_ref, s_ref = float(big.double().mean()), float(big.double().std())
    m, s = so.global_mean_std(big.to(DEV))
    assert abs(m - m_ref) < 1e-6 and abs(s - s_ref) < 1e-6


@pytest.mark.parametrize("shape,f,dims,cl", [((3, 5, 64, 64, 2), 4, 2, True), ((2, 3, 33, 47), 2, 2, False), ((4, 1, 1024), 8, 1, False),
                                              ((2, 6, 50, 1), 3, 1, True)])
def test_stride_downsample_is_the_reference_slice(shape, f, dims, cl):
    from oracle import spectral_ops_port as op
    import resolution_pde_b200.spectral_ops as so
    x = torch.randn(*shape)
    want = op.stride_downsample(x, f, dims, cl)
    got = so.stride_downsample(x.to(DEV), f, dims, cl)
    assert tuple(got.shape) == tuple(want.shape) and torch.equal(got.cpu(), want)


def test_binned_sums_kernel():
    import resolution_pde_b200 as R
    lib = R._lib.load()
    g = torch.Generator().manual_seed(1)
    p = torch.rand(2, 5000, dtype=torch.float64, generator=g).to(DEV)
    b = torch.randint(-1, 9, (5000,), generator=g, dtype=torch.int32).to(DEV)
    out = torch.empty(2, 9, dtype=torch.float64, device=DEV)
    assert lib.rpde_binned_sums_f64(p.data_ptr(), b.data_ptr(), out.data_ptr(), 2, 9, 5000, torch.cuda.current_stream().cuda_stream) == 0
    want = torch.stack([torch.stack([p[r][b == k].sum() for k in range(9)]) for r in range(2)])
    assert torch.allclose(out, want, rtol=1e-12)
