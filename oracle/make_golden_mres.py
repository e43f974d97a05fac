"""Generate tests/golden/training/mres_fno2d.npz: the UNMODIFIED reference `train.mres_training.ResolutionGroupedDataLoader`
(train/mres_training.py:75-131) feeding the UNMODIFIED `train.training.train` loop (train/training.py:19-88) on the
reference's own FNO2d, CPU, build container only.   Run:  python oracle/make_golden_mres.py   (needs /root/reference)

Dataset: a true multi-resolution mix (BASELINE config 4 at test size): 10 samples at 16x16, 5 at 24x24, 3 at 32x32,
batch size 4, shuffle=False, so the grouped loader yields batches of 4,4,2 / 4,1 / 3 samples (ragged tails included,
SURVEY section 2 row 7).  Stored: the samples, the batch sequence the reference loader produced (resolution and size of every
batch), train / validation loss histories of 4 epochs, initial and final state-dict.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from make_golden import import_reference
from make_golden_training import import_train

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "training")
EPOCHS, LR, BS = 4, 1e-2, 4
MIX = ((16, 10), (24, 5), (32, 3))


def field(n, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(1, n, n, generator=g)
    f = torch.fft.rfft2(x)
    kx = torch.fft.fftfreq(n, 1.0 / n).abs().reshape(-1, 1)
    ky = torch.arange(f.shape[-1]).reshape(1, -1)
    x = torch.fft.irfft2(f * torch.exp(-(kx ** 2 + ky ** 2) / 18.0), s=(n, n))
    x = x / x.std()
    return x, torch.roll(x, (1, 2), (-2, -1)) * 0.7 + 0.1 * x ** 2


def main():
    _, fno, _, _ = import_reference()
    tr = import_train()
    import train.mres_training as mres
    samples = []
    seed = 100
    order = []                      # interleave the resolutions in the dataset: the loader must regroup them
    for r in range(10):
        for n, cnt in MIX:
            if r < cnt:
                order.append(n)
    for n in order:
        samples.append(field(n, seed))
        seed += 1
    val = [field(16, 900 + i) for i in range(4)] + [field(32, 950 + i) for i in range(2)]
    loader = mres.ResolutionGroupedDataLoader(samples, batch_size=BS, shuffle=False)
    vloader = mres.ResolutionGroupedDataLoader(val, batch_size=BS, shuffle=False)
    seq = [(int(bx.shape[-1]), int(bx.shape[0])) for bx, _ in loader]
    torch.manual_seed(31)
    m = fno.FNO2d(1, 1, modes1=4, modes2=4, width=8)
    init = {k: v.detach().clone() for k, v in m.state_dict().items()}
    opt = torch.optim.AdamW(m.parameters(), lr=LR, weight_decay=1e-4)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=100, eta_min=1e-5)
    hist, vhist = tr.train(m, loader, vloader, opt, sch, epochs=EPOCHS, device="cpu")
    rec = {"n_samples": np.asarray(len(samples)), "n_val": np.asarray(len(val)), "batch_seq": np.asarray(seq), "batch_size": np.asarray(BS),
           "loss_history": np.asarray(hist), "val_history": np.asarray(vhist), "lr": np.asarray(LR), "epochs": np.asarray(EPOCHS)}
    for i, (x, y) in enumerate(samples):
        rec[f"sx/{i}"], rec[f"sy/{i}"] = x.numpy(), y.numpy()
    for i, (x, y) in enumerate(val):
        rec[f"vx/{i}"], rec[f"vy/{i}"] = x.numpy(), y.numpy()
    for k, v in init.items():
        rec["param/" + k] = torch.view_as_real(v).numpy() if v.is_complex() else v.numpy()
    for k, v in m.state_dict().items():
        v = v.detach()
        rec["final/" + k] = torch.view_as_real(v).numpy() if v.is_complex() else v.numpy()
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "mres_fno2d.npz"), **rec)
    print("batches per epoch:", seq)
    print("train", ["%.6f" % h for h in hist], "val", ["%.6f" % h for h in vhist])


if __name__ == "__main__":
    main()
