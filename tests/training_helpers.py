"""Shared by the training-loop parity tests: fixture loading (tests/golden/training) and model hyper-parameters."""
import os

import numpy as np
import torch
import torch.nn.functional as F

TRAIN_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "training")
KIND = {"fno1d": "FNO1d", "fno2d": "FNO2d", "ffno1d": "FFNO1D", "ffno2d": "FFNO2D"}


def load_training_fixture(name):
    z = np.load(os.path.join(TRAIN_GOLD, name), allow_pickle=False)
    hyper = {k[6:]: (z[k].item() if z[k].shape == () else z[k]) for k in z.files if k.startswith("hyper/")}
    return {
        "params": {k[6:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param/")},
        "final": {k[6:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("final/")},
        "hyper": hyper, "xs": torch.from_numpy(z["xs"]), "ys": torch.from_numpy(z["ys"]),
        "vx": torch.from_numpy(z["vx"]), "vy": torch.from_numpy(z["vy"]),
        "loss_history": z["loss_history"], "val_history": z["val_history"], "lr": float(z["lr"]), "epochs": int(z["epochs"]),
    }


def oracle_hyper(name, h):
    kind = name.split("_")[1]
    if kind.startswith("fno"):
        return KIND[kind], dict(activation={"relu": F.relu, "gelu": F.gelu}[str(h["activation"])], n_blocks=int(h["n_blocks"]))
    out = dict(n_layers=int(h["n_layers"]), n_modes=int(h["n_modes"]), n_ff_layers=int(h["n_ff_layers"]), layer_norm=bool(h["layer_norm"]))
    if kind == "ffno1d":
        out["activation"] = str(h["activation"])
    else:
        out["use_grid"] = bool(h["use_grid"])
    return KIND[kind], out
