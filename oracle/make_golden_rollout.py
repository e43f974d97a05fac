"""Generate tests/golden/rollout/rollout_*.npz from the UNMODIFIED reference (build container only).

Run:  python oracle/make_golden_rollout.py        (needs /root/reference; CPU only)

1-D cases call the reference's own `perform_rollout_1d` (utils/autoregressive_step.py:284-309) on the reference's own
FNO1d / FFNO1D modules.  The evaluation around it lives inline in `evaluate_1d_rollout_all_resolution` (:141-196, needs
hydra datasets), so its arithmetic is replayed here line by line with the reference's RelativeL2Loss: encode t=0,
roll out, decode, one loss per step, mean.  `SimpleNormalizer` is defined inside a dataloader function
(dataloaders/ks_naive_markov.py:396-406) and cannot be imported; it is restated below verbatim in behaviour.
2-D cases (the reference has no 2-D rollout) run the same reference loop body on the reference's FNO2d / FFNO2D modules.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

from make_golden import REF, import_reference

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "rollout")


class SimpleNormalizer:
    def __init__(self, mean, std, eps=1e-8):
        self.mean = float(mean)
        self.std = float(std)
        self.eps = eps

    def encode(self, x):
        return (x - self.mean) / (self.std + self.eps)

    def decode(self, x, device="cuda"):
        return x * (self.std + self.eps) + self.mean


def import_rollout():
    for name in ("hydra", "hydra.utils", "omegaconf", "seaborn", "wandb"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = types.ModuleType(name)
                if name == "hydra":
                    m.__path__ = []
                sys.modules[name] = m
    sys.modules["hydra.utils"].instantiate = getattr(sys.modules["hydra.utils"], "instantiate", None)
    sys.modules["omegaconf"].OmegaConf = getattr(sys.modules["omegaconf"], "OmegaConf", object)
    import utils.autoregressive_step as ar
    return ar


def rollout_2d(model, initial_condition, rollout_steps, x_normalizer, y_normalizer):
    """the loop body of perform_rollout_1d on (B, H, W) states"""
    predictions = []
    cur = initial_condition
    for _ in range(rollout_steps):
        nxt = model(cur.unsqueeze(1))
        if nxt.dim() == 4 and nxt.shape[1] == 1:
            nxt = nxt.squeeze(1)
        predictions.append(nxt.unsqueeze(1))
        if y_normalizer is not None and x_normalizer is not None:
            cur = x_normalizer.encode(y_normalizer.decode(nxt.unsqueeze(1), device="cpu").squeeze(1).unsqueeze(1)).squeeze(1)
        else:
            cur = nxt
    return torch.cat(predictions, dim=1)


def run_case(name, ar, loss_fn, model, traj, rollout_steps, xn, yn, hyper):
    model.eval()
    with torch.no_grad():
        T = min(rollout_steps, traj.shape[1] - 1)
        ic = traj[:, 0]
        ic_n = xn.encode(ic.unsqueeze(1)).squeeze(1) if xn is not None else ic
        if traj.dim() == 3:
            preds = ar.perform_rollout_1d(model=model, initial_condition=ic_n, rollout_steps=T, model_type="ffno1d",
                                          time_val=None, device="cpu", x_normalizer=xn, y_normalizer=yn)
        else:
            preds = rollout_2d(model, ic_n, T, xn, yn)
        if yn is not None:
            flat = preds.reshape((-1, 1) + tuple(preds.shape[2:]))
            den = yn.decode(flat, device="cpu").reshape(preds.shape)
        else:
            den = preds
        truth = traj[:, 1:T + 1]
        steps = [loss_fn(den[:, t], truth[:, t]).item() for t in range(T)]
        batch = sum(steps) / len(steps)
    rec = {"traj": traj.numpy(), "preds": preds.numpy(), "step_losses": np.asarray(steps, dtype=np.float64),
           "batch_loss": np.asarray(batch), "rollout_steps": np.asarray(rollout_steps),
           "xn": np.asarray([xn.mean, xn.std, xn.eps] if xn is not None else [np.nan] * 3),
           "yn": np.asarray([yn.mean, yn.std, yn.eps] if yn is not None else [np.nan] * 3)}
    for k, v in model.state_dict().items():
        rec["param/" + k] = v.detach().numpy()
    for k, v in hyper.items():
        rec["hyper/" + k] = np.asarray(v)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(f"{name}: preds {tuple(preds.shape)} steps {['%.4f' % s for s in steps]} batch {batch:.6f}")


def main():
    sc, fno, ffno, loss_mod = import_reference()
    ar = import_rollout()
    L = loss_mod.RelativeL2Loss(size_average=True, reduction=True)
    rn = lambda *s, seed=0: torch.randn(*s, generator=torch.Generator().manual_seed(seed))
    xn, yn = SimpleNormalizer(0.31, 1.7), SimpleNormalizer(-0.22, 2.1)

    torch.manual_seed(14)
    m = fno.FNO1d(1, 1, modes=8, width=16)
    run_case("rollout_fno1d_n64_norm", ar, L, m, 0.31 + 1.7 * rn(3, 7, 64, seed=1), 5, xn, yn,
             dict(modes=8, width=16, n_blocks=4, activation="relu"))
    run_case("fno1d_n48_nonorm_clamped", ar, L, m, rn(2, 4, 48, seed=2), 10, None, None,
             dict(modes=8, width=16, n_blocks=4, activation="relu"))
    os.replace(os.path.join(OUT, "fno1d_n48_nonorm_clamped.npz"), os.path.join(OUT, "rollout_fno1d_n48_nonorm_clamped.npz"))
    torch.manual_seed(16)
    m = ffno.FFNO1D(1, 1, width=16, n_layers=2, n_modes=8, factor=4, ff_weight_norm=True, n_ff_layers=3,
                    layer_norm=True, dropout=0.0, mode="full", activation="gelu", use_grid=True)
    hy = dict(width=16, n_layers=2, n_modes=8, n_ff_layers=3, layer_norm=1, activation="gelu", ff_weight_norm=1)
    run_case("rollout_ffno1d_n32_norm", ar, L, m, 0.31 + 1.7 * rn(4, 9, 32, seed=3), 8, xn, yn, hy)
    run_case("rollout_ffno1d_n33_ynorm_only", ar, L, m, rn(2, 5, 33, seed=4), 4, None, yn, hy)
    torch.manual_seed(15)
    m = fno.FNO2d(1, 1, modes1=4, modes2=4, width=8)
    run_case("rollout_fno2d_16x20_norm", ar, L, m, 0.31 + 1.7 * rn(2, 5, 16, 20, seed=5), 4, xn, yn,
             dict(modes1=4, modes2=4, width=8, n_blocks=4, activation="gelu"))
    torch.manual_seed(17)
    m = ffno.FFNO2D(1, 1, width=16, n_layers=2, n_modes=6, factor=4, ff_weight_norm=True, n_ff_layers=3,
                    layer_norm=True, dropout=0.0, mode="full", use_grid=True)
    run_case("rollout_ffno2d_16x12_norm", ar, L, m, 0.31 + 1.7 * rn(2, 6, 16, 12, seed=6), 5, xn, yn,
             dict(width=16, n_layers=2, n_modes=6, n_ff_layers=3, layer_norm=1, ff_weight_norm=1, use_grid=1))


if __name__ == "__main__":
    main()
