"""Generate tests/golden/*.npz from the UNMODIFIED reference (build container only).

Run:  python oracle/make_golden.py            (needs /root/reference; CPU only)

The reference modules import h5py / matplotlib at module top but never use them on
this path, so empty stubs are injected.  For every case we store the
seeded input, the target, the reference state-dict, the reference forward output,
the RelativeL2 loss and the gradients of the loss w.r.t. every parameter and the
input.  Nothing under tests/ or the product reads /root/reference at run time; the
fixtures are the pin.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("RPDE_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def import_reference():
    sys.dont_write_bytecode = True
    for name in ("h5py", "matplotlib", "matplotlib.pyplot", "matplotlib.patches"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            if name == "matplotlib":
                m.__path__ = []
            sys.modules[name] = m
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import models.spectral_convolution as sc
    import models.fno as fno
    import models.ffno as ffno
    import utils.loss as loss
    return sc, fno, ffno, loss


def _np(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().numpy()


def run_case(name, module, x, loss_fn, hyper, out_index=None):
    """forward + rel-L2 against a seeded target + backward; dump everything."""
    x = x.clone().requires_grad_(True)
    out = module(x)
    if out_index is not None:
        out = out[out_index]
    g = torch.Generator().manual_seed(4321)
    y = out.detach() + 0.5 * torch.randn(out.shape, generator=g)
    loss = loss_fn(out, y)
    loss.backward()
    rec = {"x": _np(x), "y": _np(y), "out": _np(out), "loss": _np(loss), "grad_x": _np(x.grad)}
    for k, v in module.state_dict().items():
        rec["param/" + k] = _np(v)
    for k, p in module.named_parameters():
        rec["grad/" + k] = _np(p.grad if p.grad is not None else torch.zeros_like(p))
    for k, v in hyper.items():
        rec["hyper/" + k] = np.asarray(v)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(f"{name}: out {tuple(out.shape)} loss {float(loss):.6f}")


def main():
    sc, fno, ffno, loss_mod = import_reference()
    L = loss_mod.RelativeL2Loss()
    rn = lambda *s, seed=0: torch.randn(*s, generator=torch.Generator().manual_seed(seed))

    # ---- FNO spectral layers -------------------------------------------------
    for (N, modes) in ((32, 5), (33, 7), (16, 9)):       # 16/9 keeps the Nyquist bin
        torch.manual_seed(0)
        m = sc.SpectralConv1d(3, 4, modes)
        with torch.no_grad():
            m.weights1.mul_(12.0)                          # O(1) outputs
        run_case(f"spectralconv1d_n{N}_k{modes}", m, rn(2, 3, N, seed=N), L, dict(modes=modes))
    for (H, W, m1, m2) in ((16, 12, 3, 4), (15, 9, 2, 5), (8, 8, 4, 5)):
        torch.manual_seed(1)
        m = sc.SpectralConv2d(3, 2, m1, m2)
        with torch.no_grad():
            m.weights1.mul_(6.0)
            m.weights2.mul_(6.0)
        run_case(f"spectralconv2d_{H}x{W}_m{m1}_{m2}", m, rn(2, 3, H, W, seed=H * W), L, dict(m1=m1, m2=m2))

    # ---- F-FNO spectral layers ------------------------------------------------
    for (N, modes, mode, norm, act) in ((32, 6, "full", "ortho", "identity"), (33, 40, "full", "ortho", "gelu"),
                                        (32, 6, "low-pass", "ortho", "identity"), (24, 5, "full", "backward", "relu")):
        torch.manual_seed(2)
        m = sc.FSpectralConv1d(8, modes, factor=2, n_ff_layers=3, layer_norm=True, mode=mode, fft_norm=norm,
                               activation=act)
        run_case(f"fspectralconv1d_n{N}_k{modes}_{mode}_{norm}_{act}", m, rn(2, N, 8, seed=N + modes), L,
                 dict(modes=modes, n_ff_layers=3, layer_norm=1, mode=mode, fft_norm=norm, activation=act), out_index=0)
    for (M, N, modes, mode, ln) in ((16, 12, 5, "full", True), (9, 10, 16, "full", False), (8, 8, 3, "low-pass", True)):
        torch.manual_seed(3)
        m = sc.FSpectralConv2d(8, modes, factor=2, n_ff_layers=2, layer_norm=ln, mode=mode)
        run_case(f"fspectralconv2d_{M}x{N}_k{modes}_{mode}_ln{int(ln)}", m, rn(2, M, N, 8, seed=M * N + modes), L,
                 dict(modes=modes, n_ff_layers=2, layer_norm=int(ln), mode=mode), out_index=0)

    # ---- whole models ----------------------------------------------------------
    torch.manual_seed(4)
    m = fno.FNO1d(1, 1, modes=8, width=16)
    run_case("fno1d_n64", m, rn(2, 1, 64, seed=64), L, dict(modes=8, width=16, n_blocks=4, activation="relu"))
    torch.manual_seed(5)
    m = fno.FNO2d(1, 1, modes1=4, modes2=4, width=8)
    run_case("fno2d_16x20", m, rn(2, 1, 16, 20, seed=320), L, dict(modes1=4, modes2=4, width=8, n_blocks=4, activation="gelu"))
    torch.manual_seed(6)
    m = ffno.FFNO1D(1, 1, width=16, n_layers=2, n_modes=8, factor=4, ff_weight_norm=True, n_ff_layers=3,
                    layer_norm=True, dropout=0.0, mode="full", activation="gelu", use_grid=True)
    run_case("ffno1d_n32", m, rn(2, 1, 32, seed=32), L,
             dict(width=16, n_layers=2, n_modes=8, n_ff_layers=3, layer_norm=1, activation="gelu", ff_weight_norm=1))
    torch.manual_seed(7)
    m = ffno.FFNO2D(1, 1, width=16, n_layers=2, n_modes=6, factor=4, ff_weight_norm=True, n_ff_layers=3,
                    layer_norm=True, dropout=0.0, mode="full", use_grid=True)
    run_case("ffno2d_16x12", m, rn(2, 1, 16, 12, seed=192), L,
             dict(width=16, n_layers=2, n_modes=6, n_ff_layers=3, layer_norm=1, ff_weight_norm=1, use_grid=1))
    torch.manual_seed(8)
    m = ffno.FFNO2D(2, 3, width=8, n_layers=1, n_modes=4, factor=2, ff_weight_norm=False, n_ff_layers=2,
                    layer_norm=False, dropout=0.0, mode="full", use_grid=False)
    run_case("ffno2d_nogrid_8x8", m, rn(2, 2, 8, 8, seed=88), L,
             dict(width=8, n_layers=1, n_modes=4, n_ff_layers=2, layer_norm=0, ff_weight_norm=0, use_grid=0))

    # ---- loss -------------------------------------------------------------------
    a, b = rn(5, 3, 17, seed=11), rn(5, 3, 17, seed=12)
    rec = {"x": _np(a), "y": _np(b)}
    for sa, red, tag in ((True, True, "mean"), (False, True, "sum"), (True, False, "none")):
        rec[tag] = _np(loss_mod.RelativeL2Loss(size_average=sa, reduction=red)(a, b))
    np.savez_compressed(os.path.join(OUT, "relative_l2.npz"), **rec)
    print("relative_l2 done")


if __name__ == "__main__":
    main()
