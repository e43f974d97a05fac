"""Shared by the rollout tests: fixture loading and the oracle's model forward for a fixture."""
import os

import numpy as np
import torch
import torch.nn.functional as F

ROLLOUT_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rollout")


def _norm(a):
    return None if np.isnan(a[0]) else (float(a[0]), float(a[1]), float(a[2]))


def load_rollout_fixture(name):
    z = np.load(os.path.join(ROLLOUT_GOLD, name), allow_pickle=False)
    return {
        "params": {k[6:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param/")},
        "hyper": {k[6:]: (z[k].item() if z[k].shape == () else z[k]) for k in z.files if k.startswith("hyper/")},
        "traj": torch.from_numpy(z["traj"]), "preds": torch.from_numpy(z["preds"]),
        "step_losses": z["step_losses"], "batch_loss": float(z["batch_loss"]), "rollout_steps": int(z["rollout_steps"]),
        "xn": _norm(z["xn"]), "yn": _norm(z["yn"]),
    }


def model_kind(name):
    return name.split("_")[1]          # rollout_<kind>_...


def oracle_forward(name, params, hyper):
    """x -> model(x) through the CPU oracle's functional model ports"""
    from oracle import reference_port as rp
    h, kind = hyper, model_kind(name)
    act = {"relu": F.relu, "gelu": F.gelu}
    if kind == "fno1d":
        return lambda x: rp.fno1d_forward(params, x, activation=act[str(h["activation"])], n_blocks=int(h["n_blocks"]))
    if kind == "fno2d":
        return lambda x: rp.fno2d_forward(params, x, activation=act[str(h["activation"])], n_blocks=int(h["n_blocks"]))
    if kind == "ffno1d":
        return lambda x: rp.ffno1d_forward(params, x, n_layers=int(h["n_layers"]), n_modes=int(h["n_modes"]),
                                           n_ff_layers=int(h["n_ff_layers"]), layer_norm=bool(h["layer_norm"]),
                                           activation=str(h["activation"]))
    if kind == "ffno2d":
        return lambda x: rp.ffno2d_forward(params, x, n_layers=int(h["n_layers"]), n_modes=int(h["n_modes"]),
                                           n_ff_layers=int(h["n_ff_layers"]), layer_norm=bool(h["layer_norm"]),
                                           use_grid=bool(h["use_grid"]))
    raise KeyError(name)


class Norm:
    """(mean, std, eps) tuple -> the attribute form the product's rollout accepts"""

    def __init__(self, t):
        self.mean, self.std, self.eps = t
