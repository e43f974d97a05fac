"""Generate tests/golden/training/train_*.npz: the UNMODIFIED reference `train.training.train` loop (train/training.py:19-88)
run on the reference's own modules for 5 epochs x 4 batches = 20 optimizer steps on the CPU (build container only).

Run:  python oracle/make_golden_training.py        (needs /root/reference)
Optimizer and scheduler as main_1d.py:144-145 builds them: AdamW(lr, weight_decay=1e-4), CosineAnnealingLR(T_max=100,
eta_min=1e-5), stepped once per epoch (training.py:76-79).  `wandb` is an empty stub whose log() does nothing.
Stored: initial state-dict, the batches, per-epoch train / validation loss histories, final state-dict.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

from make_golden import import_reference

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "training")
EPOCHS, NB, LR = 5, 4, 1e-3


def import_train():
    if "wandb" not in sys.modules:
        w = types.ModuleType("wandb")
        w.log = lambda *a, **k: None
        sys.modules["wandb"] = w
    import train.training as tr
    return tr


def run_case(name, tr, model, xs, ys, vx, vy, hyper):
    init = {k: v.detach().clone() for k, v in model.state_dict().items()}
    opt = torch.optim.AdamW(model.parameters(), lr=LR, weight_decay=1e-4)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=100, eta_min=1e-5)
    hist, vhist = tr.train(model, list(zip(xs, ys)), [(vx, vy)], opt, sch, epochs=EPOCHS, device="cpu")
    rec = {"xs": torch.stack(xs).numpy(), "ys": torch.stack(ys).numpy(), "vx": vx.numpy(), "vy": vy.numpy(),
           "loss_history": np.asarray(hist), "val_history": np.asarray(vhist), "lr": np.asarray(LR), "epochs": np.asarray(EPOCHS)}
    for k, v in init.items():
        rec["param/" + k] = v.numpy()
    for k, v in model.state_dict().items():
        rec["final/" + k] = v.detach().numpy()
    for k, v in hyper.items():
        rec["hyper/" + k] = np.asarray(v)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(name, ["%.5f" % h for h in hist], ["%.5f" % h for h in vhist])


def smooth(shape, seed):
    """band-limited fields and a target that is a smooth functional of them (so the loss really decreases)"""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(*shape, generator=g)
    f = torch.fft.rfft(x, dim=-1)
    k = torch.arange(f.shape[-1])
    x = torch.fft.irfft(f * torch.exp(-(k / 6.0) ** 2), n=shape[-1], dim=-1)
    x = x / x.std()
    return x, torch.roll(x, 1, -1) * 0.8 + 0.1 * x ** 2


def main():
    sc, fno, ffno, _ = import_reference()
    tr = import_train()
    torch.manual_seed(21)
    m = fno.FNO1d(1, 1, modes=8, width=16)
    data = [smooth((4, 1, 64), s) for s in range(NB + 1)]
    run_case("train_fno1d_n64", tr, m, [d[0] for d in data[:NB]], [d[1] for d in data[:NB]], data[NB][0], data[NB][1],
             dict(modes=8, width=16, n_blocks=4, activation="relu"))
    torch.manual_seed(22)
    m = ffno.FFNO2D(1, 1, width=16, n_layers=2, n_modes=6, factor=4, ff_weight_norm=True, n_ff_layers=3,
                    layer_norm=True, dropout=0.0, mode="full", use_grid=True)
    data = [smooth((3, 1, 16, 12), 10 + s) for s in range(NB + 1)]
    run_case("train_ffno2d_16x12", tr, m, [d[0] for d in data[:NB]], [d[1] for d in data[:NB]], data[NB][0], data[NB][1],
             dict(width=16, n_layers=2, n_modes=6, n_ff_layers=3, layer_norm=1, ff_weight_norm=1, use_grid=1))
    torch.manual_seed(23)
    m = ffno.FFNO1D(1, 1, width=16, n_layers=2, n_modes=8, factor=4, ff_weight_norm=True, n_ff_layers=3,
                    layer_norm=True, dropout=0.0, mode="full", activation="gelu", use_grid=True)
    data = [smooth((4, 1, 32), 20 + s) for s in range(NB + 1)]
    run_case("train_ffno1d_n32", tr, m, [d[0] for d in data[:NB]], [d[1] for d in data[:NB]], data[NB][0], data[NB][1],
             dict(width=16, n_layers=2, n_modes=8, n_ff_layers=3, layer_norm=1, activation="gelu", ff_weight_norm=1))


if __name__ == "__main__":
    main()
