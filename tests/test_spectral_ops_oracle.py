"""CPU: the restatement of the reference's spectral data-preparation operators (oracle/spectral_ops_port.py) against
the fixtures generated from the unmodified reference (oracle/make_golden_spectral_ops.py)."""
import glob
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "specops")
CASES = sorted(glob.glob(os.path.join(GOLDEN, "specops_*.npz")))


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
    assert len(CASES) == 22


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[8:-4] for p in CASES])
def test_oracle_matches_reference_fixture(path):
    from oracle import spectral_ops_port as port
    y, ref = run_case(port, path)
    assert y.shape == ref.shape
    assert rel(y, ref) < 2e-6
