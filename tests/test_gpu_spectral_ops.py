"""GPU: resolution_pde_b200.spectral_ops (truncated-DFT contractions through the C ABI) against the reference fixtures
and against the fp64 oracle on seeded inputs at dataset sizes.  Tolerance 1e-5 relative L2 (fp32 / 3xTF32 engines)."""
import os

import pytest
import torch

from test_spectral_ops_oracle import CASES, DOWN_CASES, FREQ_CASES, rel, run_case, run_down_case, run_freq_case, vec_rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-5


class _OnDevice:
    """spectral_ops with the fixture's CPU input moved to the device"""

    def __init__(self):
        import resolution_pde_b200.spectral_ops as so
        self.so = so

    def __getattr__(self, name):
        fn = getattr(self.so, name)
        return lambda x, *a, **k: fn(x.to(DEV), *a, **k)


@pytest.fixture(params=["fp32", "3xtf32"])
def engine(request):
    import resolution_pde_b200 as R
    R.set_precision(request.param)
    yield request.param
    R.set_precision("fp32")


@pytest.mark.parametrize("path", CASES, ids=lambda p: os.path.basename(p)[8:-4])
def test_cuda_matches_reference_fixture(engine, path):
    y, ref = run_case(_OnDevice(), path)
    assert tuple(y.shape) == tuple(ref.shape)
    assert rel(y, ref) < TOL


def test_cpu_tensors_are_refused():
    import resolution_pde_b200.spectral_ops as so
    with pytest.raises(ValueError):
        so.lowpass_filter_1d(torch.randn(2, 3, 16), 0.5)            # CPU tensor: no fallback


@pytest.mark.parametrize("case", ["lp1d", "lp2d", "rs1d_down", "rs1d_up", "rs2d_down", "rs2d_up"])
def test_cuda_matches_fp64_oracle_at_dataset_sizes(engine, case):
    from oracle import spectral_ops_port as port
    import resolution_pde_b200.spectral_ops as so
    torch.manual_seed(7)
    if case == "lp1d":
        x = torch.randn(8, 5, 2, 1024); y = so.lowpass_filter_1d(x.to(DEV), 0.25); ref = port.lowpass_filter_1d(x.double(), 0.25)
    elif case == "lp2d":
        x = torch.randn(2, 3, 1, 256, 256); y = so.lowpass_filter_2d(x.to(DEV), 0.125); ref = port.lowpass_filter_2d(x.double(), 0.125)
    elif case == "rs1d_down":
        x = torch.randn(16, 4, 2048); y = so.resize_1d(x.to(DEV), 256); ref = port.resize_1d(x.double(), 256)
    elif case == "rs1d_up":
        x = torch.randn(16, 4, 128); y = so.resize_1d(x.to(DEV), 1024); ref = port.resize_1d(x.double(), 1024)
    elif case == "rs2d_down":
        x = torch.randn(3, 2, 256, 256); y = so.resize(x.to(DEV), (64, 64)); ref = port.resize(x.double(), (64, 64))
    else:
        x = torch.randn(3, 2, 64, 64); y = so.resize(x.to(DEV), (256, 256)); ref = port.resize(x.double(), (256, 256))
    assert rel(y, ref) < TOL


@pytest.mark.parametrize("path", FREQ_CASES, ids=lambda p: os.path.basename(p)[8:-4])
def test_frequency_error_matches_reference_fixture(engine, path):
    """Parseval route on the device against the reference's mode-by-mode inverse FFTs."""
    import resolution_pde_b200.frequency_error as fe
    (e, m, f), (re, rm, rf) = run_freq_case(fe, path, to=DEV)
    assert e.shape == re.shape and m.shape == rm.shape and f.shape == rf.shape
    assert vec_rel(e, re) < TOL and vec_rel(m, rm) < TOL and vec_rel(f, rf) < 1e-7
    assert abs(e - re).max() < 1e-5 * abs(re).max() and abs(m - rm).max() < 1e-5 * abs(rm).max()


def test_frequency_error_at_dataset_sizes(engine):
    from oracle import spectral_ops_port as port
    import resolution_pde_b200.frequency_error as fe
    torch.manual_seed(11)
    yh, yt = torch.randn(8, 1, 1024), torch.randn(8, 1, 1024)
    e, m, f = fe.decompose_error_by_frequency_1d(yh.to(DEV), yt.to(DEV), 64)
    re, rm, rf = port.decompose_error_by_frequency_1d(yh.double(), yt.double(), 64)
    assert vec_rel(e, re) < TOL and vec_rel(m, rm) < TOL
    yh, yt = torch.randn(2, 1, 128, 128), torch.randn(2, 1, 128, 128)
    e, m, f = fe.decompose_error_by_frequency_2d(yh.to(DEV), yt.to(DEV), 32)
    re, rm, rf = port.decompose_error_by_frequency_2d(yh.double(), yt.double(), 32)
    assert vec_rel(e, re) < TOL and vec_rel(m, rm) < TOL and vec_rel(f, rf) < 1e-12


@pytest.mark.parametrize("path", DOWN_CASES, ids=lambda p: os.path.basename(p)[8:-4])
def test_downsample_cuda_matches_reference_fixture(engine, path):
    import resolution_pde_b200.spectral_ops as so
    y, ref = run_down_case(so, path, to=DEV)
    assert tuple(y.shape) == tuple(ref.shape)
    assert rel(y, ref) < TOL


@pytest.mark.parametrize("shape,n", [((4, 2, 256, 256), 64), ((2, 1, 512, 512), 128), ((16, 3, 2048), 256), ((8, 1024), 33)])
def test_downsample_cuda_matches_fp64_oracle_at_dataset_sizes(engine, shape, n):
    from oracle import spectral_ops_port as port
    import resolution_pde_b200.spectral_ops as so
    x = torch.randn(*shape, generator=torch.Generator().manual_seed(n))
    if len(shape) == 4:
        y, ref = so.downsample(x.to(DEV), n), port.downsample(x.double(), n)
    else:
        y, ref = so.downsample_1d(x.to(DEV), n), port.downsample_1d(x.double(), n)
    assert tuple(y.shape) == tuple(ref.shape)
    assert rel(y, ref) < TOL
