"""CPU: the restatement of the reference's spectral data-preparation operators (oracle/spectral_ops_port.py) against
the fixtures generated from the unmodified reference (oracle/make_golden_spectral_ops.py)."""
import glob
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "specops")
CASES = sorted(p for p in glob.glob(os.path.join(GOLDEN, "specops_*.npz")) if "freqerr" not in p and "downsample" not in p)
DOWN_CASES = sorted(glob.glob(os.path.join(GOLDEN, "specops_downsample*.npz")))
FREQ_CASES = sorted(glob.glob(os.path.join(GOLDEN, "specops_freqerr*.npz")))


def run_case(mod, path, dtype=None):
    d = np.load(path)
    x = torch.from_numpy(d["x"])
    x = x.to(dtype) if dtype is not None else x
    name = os.path.basename(path)
    if name.startswith("specops_lowpass1d"):
        y = mod.lowpass_filter_1d(x, float(d["ratio"]))
    elif name.startswith("specops_lowpass2d"):
        y = mod.lowpass_filter_2d(x, float(d["ratio"]))
    elif name.startswith("specops_resize1d"):
        y = mod.resize_1d(x, int(d["out"]))
    else:
        y = mod.resize(x, tuple(int(v) for v in d["out"]), permute=bool(d["permute"]))
    return y, torch.from_numpy(d["y"])


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def test_fixtures_present():
    assert len(CASES) == 26 and len(FREQ_CASES) == 8 and len(DOWN_CASES) == 13


def run_freq_case(mod, path, to=None):
    d = np.load(path)
    yh, yt = torch.from_numpy(d["y_hat"]), torch.from_numpy(d["x"])
    if to is not None:
        yh, yt = yh.to(to), yt.to(to)
    num = int(d["num"])
    if "freqerr1d" in path:
        got = mod.decompose_error_by_frequency_1d(yh, yt, None if num < 0 else num)
    else:
        got = mod.decompose_error_by_frequency_2d(yh, yt, num)
    return got, (d["err"], d["mag"], d["freqs"])


def vec_rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


@pytest.mark.parametrize("path", FREQ_CASES, ids=[os.path.basename(p)[8:-4] for p in FREQ_CASES])
def test_frequency_error_oracle_matches_reference_fixture(path):
    from oracle import spectral_ops_port as port
    (e, m, f), (re, rm, rf) = run_freq_case(port, path)
    assert e.shape == re.shape and m.shape == rm.shape and f.shape == rf.shape
    assert vec_rel(e, re) < 2e-6 and vec_rel(m, rm) < 2e-6 and vec_rel(f, rf) < 1e-7


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[8:-4] for p in CASES])
def test_oracle_matches_reference_fixture(path):
    from oracle import spectral_ops_port as port
    y, ref = run_case(port, path)
    assert y.shape == ref.shape
    assert rel(y, ref) < 2e-6


def run_down_case(mod, path, to=None, dtype=None):
    """downsample_1d / downsample (the reference runs scipy on numpy arrays; both ports take tensors)"""
    d = np.load(path)
    x = torch.from_numpy(d["x"])
    x = x.to(dtype) if dtype is not None else x
    x = x.to(to) if to is not None else x
    fn = mod.downsample_1d if "downsample1d" in path else mod.downsample
    return fn(x, int(d["out"])), torch.from_numpy(d["y"])


@pytest.mark.parametrize("path", DOWN_CASES, ids=[os.path.basename(p)[8:-4] for p in DOWN_CASES])
def test_downsample_oracle_matches_reference_fixture(path):
    from oracle import spectral_ops_port as port
    y, ref = run_down_case(port, path)
    assert y.shape == ref.shape
    assert rel(y, ref) < 2e-6
