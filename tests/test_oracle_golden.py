"""Pins the CPU oracle (oracle/reference_port.py) against fixtures generated from the UNMODIFIED
reference (oracle/make_golden.py, run in the build container).  CPU only."""
import glob
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import reference_port as rp

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-5  # relative L2 per tensor, fp32 (BASELINE.json north_star)


def rel(a, b):
    a = torch.as_tensor(a).to(torch.complex128 if torch.as_tensor(a).is_complex() else torch.float64)
    b = torch.as_tensor(b).to(a.dtype)
    den = b.abs().pow(2).sum().sqrt().item()
    num = (a - b).abs().pow(2).sum().sqrt().item()
    return num / den if den > 0 else num


def load(name):
    z = np.load(os.path.join(GOLD, name), allow_pickle=False)
    params = {k[6:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param/")}
    grads = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("grad/")}
    hyper = {k[6:]: z[k] for k in z.files if k.startswith("hyper/")}
    return z, params, grads, hyper


def run(fn, params, x, y):
    ps = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    xx = x.clone().requires_grad_(True)
    out = fn(ps, xx)
    loss = rp.relative_l2(out, y)
    names = list(ps)
    gs = torch.autograd.grad(loss, [xx] + [ps[k] for k in names], allow_unused=True)
    g = {k: (v if v is not None else torch.zeros_like(ps[k])) for k, v in zip(names, gs[1:])}
    return out.detach(), loss.detach(), gs[0], g


def case_fn(name, hyper):
    h = {k: (v.item() if v.shape == () else v) for k, v in hyper.items()}
    if name.startswith("spectralconv1d"):
        return lambda p, x: rp.spectral_conv1d(x, p["weights1"])
    if name.startswith("spectralconv2d"):
        return lambda p, x: rp.spectral_conv2d(x, p["weights1"], p["weights2"])
    if name.startswith("fspectralconv1d"):
        return lambda p, x: rp.fspectral_conv1d(x, p, int(h["modes"]), int(h["n_ff_layers"]), bool(h["layer_norm"]),
                                                str(h["mode"]), str(h["fft_norm"]), str(h["activation"]))[0]
    if name.startswith("fspectralconv2d"):
        return lambda p, x: rp.fspectral_conv2d(x, p, int(h["modes"]), int(h["n_ff_layers"]), bool(h["layer_norm"]),
                                                str(h["mode"]))[0]
    if name.startswith("fno1d"):
        return lambda p, x: rp.fno1d_forward(p, x, activation=str(h["activation"]), n_blocks=int(h["n_blocks"]))
    if name.startswith("fno2d"):
        return lambda p, x: rp.fno2d_forward(p, x, activation=str(h["activation"]), n_blocks=int(h["n_blocks"]))
    if name.startswith("ffno1d"):
        return lambda p, x: rp.ffno1d_forward(p, x, n_layers=int(h["n_layers"]), n_modes=int(h["n_modes"]),
                                              n_ff_layers=int(h["n_ff_layers"]), layer_norm=bool(h["layer_norm"]),
                                              activation=str(h["activation"]))
    if name.startswith("ffno2d"):
        return lambda p, x: rp.ffno2d_forward(p, x, n_layers=int(h["n_layers"]), n_modes=int(h["n_modes"]),
                                              n_ff_layers=int(h["n_ff_layers"]), layer_norm=bool(h["layer_norm"]),
                                              use_grid=bool(h["use_grid"]))
    raise KeyError(name)


CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*.npz")) if "relative_l2" not in p)


def test_fixtures_present():
    assert len(CASES) >= 18


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_fixture(name):
    z, params, grads, hyper = load(name)
    x, y = torch.from_numpy(z["x"]), torch.from_numpy(z["y"])
    out, loss, gx, g = run(case_fn(name, hyper), params, x, y)
    assert rel(out, z["out"]) < TOL
    assert abs(loss.item() - float(z["loss"])) < TOL * max(1.0, abs(float(z["loss"])))
    assert rel(gx, z["grad_x"]) < TOL
    for k, ref in grads.items():
        if ref.abs().max() == 0:
            assert g[k].abs().max() == 0, k
        elif ref.abs().max() < 1e-8:
            # analytically-zero gradient (weight_v of a 1-column weight-normed Linear): round-off only
            assert g[k].abs().max() < 1e-8, k
        else:
            assert rel(g[k], ref) < TOL, k


def test_relative_l2_fixture():
    z = np.load(os.path.join(GOLD, "relative_l2.npz"))
    a, b = torch.from_numpy(z["x"]), torch.from_numpy(z["y"])
    assert rel(rp.relative_l2(a, b, True, True), z["mean"]) < 1e-6
    assert rel(rp.relative_l2(a, b, False, True), z["sum"]) < 1e-6
    assert rel(rp.relative_l2(a, b, True, False), z["none"]) < 1e-6


def test_fno_no_clamp_raises():
    w = torch.randn(2, 2, 20, dtype=torch.cfloat)
    with pytest.raises(RuntimeError):
        rp.spectral_conv1d(torch.randn(1, 2, 32), w)


def test_ffno1d_unknown_mode_raises():
    with pytest.raises(ValueError):
        rp.fspectral_fourier1d(torch.randn(1, 16, 4), torch.randn(4, 4, 3, 2), 3, mode="bogus")
