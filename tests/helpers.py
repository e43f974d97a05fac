"""Shared helpers for the parity tests: build a candidate module for a golden fixture, compare tensors."""
import os

import numpy as np
import torch
import torch.nn.functional as F

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-5  # relative L2 per tensor, fp32 engine (BASELINE.json north_star)


def rel(a, b):
    a = torch.as_tensor(a).detach().cpu()
    b = torch.as_tensor(b).detach().cpu()
    a = a.to(torch.complex128 if a.is_complex() else torch.float64)
    b = b.to(a.dtype)
    den = b.abs().pow(2).sum().sqrt().item()
    num = (a - b).abs().pow(2).sum().sqrt().item()
    return num / den if den > 0 else num


def load_fixture(name):
    z = np.load(os.path.join(GOLD, name), allow_pickle=False)
    params = {k[6:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param/")}
    grads = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("grad/")}
    hyper = {k[6:]: (z[k].item() if z[k].shape == () else z[k]) for k in z.files if k.startswith("hyper/")}
    return z, params, grads, hyper


def build_candidate(name, params, hyper):
    """Instantiate the drop-in module for a fixture; returns (module, out_index)."""
    import resolution_pde_b200 as R
    h = hyper
    if name.startswith("spectralconv1d"):
        i, o, k = params["weights1"].shape
        return R.SpectralConv1d(i, o, k), None
    if name.startswith("spectralconv2d"):
        i, o, m1, m2 = params["weights1"].shape
        return R.SpectralConv2d(i, o, m1, m2), None
    if name.startswith("fspectralconv1d"):
        c = params["fourier_weight.0"].shape[0]
        f = params["backcast_ff.layers.0.0.weight"].shape[0] // c
        return R.FSpectralConv1d(c, int(h["modes"]), factor=f, n_ff_layers=int(h["n_ff_layers"]),
                                 layer_norm=bool(h["layer_norm"]), mode=str(h["mode"]), fft_norm=str(h["fft_norm"]),
                                 activation=str(h["activation"])), 0
    if name.startswith("fspectralconv2d"):
        c = params["fourier_weight.0"].shape[0]
        f = params["backcast_ff.layers.0.0.weight"].shape[0] // c
        return R.FSpectralConv2d(c, int(h["modes"]), factor=f, n_ff_layers=int(h["n_ff_layers"]),
                                 layer_norm=bool(h["layer_norm"]), mode=str(h["mode"])), 0
    act = {"relu": F.relu, "gelu": F.gelu}
    if name.startswith("fno1d"):
        return R.FNO1d(1, 1, modes=int(h["modes"]), width=int(h["width"]), activation=act[str(h["activation"])],
                       n_blocks=int(h["n_blocks"])), None
    if name.startswith("fno2d"):
        return R.FNO2d(1, 1, modes1=int(h["modes1"]), modes2=int(h["modes2"]), width=int(h["width"]),
                       activation=act[str(h["activation"])], n_blocks=int(h["n_blocks"])), None
    if name.startswith("ffno1d"):
        return R.FFNO1D(1, 1, width=int(h["width"]), n_layers=int(h["n_layers"]), n_modes=int(h["n_modes"]), factor=4,
                        ff_weight_norm=bool(h["ff_weight_norm"]), n_ff_layers=int(h["n_ff_layers"]),
                        layer_norm=bool(h["layer_norm"]), dropout=0.0, mode="full", activation=str(h["activation"]),
                        use_grid=True), None
    if name.startswith("ffno2d"):
        inc = params["in_proj.bias"].shape[0]
        wkey = "in_proj.weight_v" if "in_proj.weight_v" in params else "in_proj.weight"
        in_lin = params[wkey].shape[1]
        use_grid = bool(h["use_grid"])
        in_ch = in_lin - 2 if use_grid else in_lin
        okey = "out_proj.weight_v" if "out_proj.weight_v" in params else "out_proj.weight"
        out_ch = params[okey].shape[0]
        f = params["fourier_layers.0.backcast_ff.layers.0.0.weight"].shape[0] // inc
        return R.FFNO2D(in_ch, out_ch, width=int(h["width"]), n_layers=int(h["n_layers"]), n_modes=int(h["n_modes"]),
                        factor=f, ff_weight_norm=bool(h["ff_weight_norm"]), n_ff_layers=int(h["n_ff_layers"]),
                        layer_norm=bool(h["layer_norm"]), dropout=0.0, mode="full", use_grid=use_grid), None
    raise KeyError(name)


def assert_grads_close(named_grads, ref_grads, tol=TOL):
    for k, ref in ref_grads.items():
        g = named_grads[k]
        if ref.abs().max() == 0:
            assert g.abs().max().item() == 0, k
        elif ref.abs().max() < 1e-8:
            assert g.abs().max().item() < 1e-7, k  # analytically zero (weight_v of a 1-column weight-normed Linear)
        else:
            assert rel(g, ref) < tol, (k, rel(g, ref))
