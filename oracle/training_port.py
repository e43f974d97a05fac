"""CPU restatement of the reference's training loop -- TEST INFRASTRUCTURE (only tests/ may import it).

Follows `train/training.py:19-88` (`train`): per epoch, for every batch zero_grad -> forward -> RelativeL2Loss -> backward ->
optimizer.step, the epoch's mean training loss, then the mean validation loss under no_grad and one scheduler step; with the
optimizer and scheduler `main_1d.py:144-145` builds (AdamW(lr, weight_decay=1e-4), CosineAnnealingLR(T_max=100, eta_min=1e-5)).
The model is the functional port of oracle/reference_port.py over a dict of leaf tensors.
Pinned by tests/golden/training/*.npz, written by the UNMODIFIED reference `train()` (oracle/make_golden_training.py).
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch

from . import reference_port as rp


def train(kind: str, params: Dict[str, torch.Tensor], batches: Sequence[Tuple[torch.Tensor, torch.Tensor]],
          val: Sequence[Tuple[torch.Tensor, torch.Tensor]], lr: float, epochs: int, **hyper):
    ps = {k: v.detach().clone().requires_grad_(v.is_floating_point() or v.is_complex()) for k, v in params.items()}
    opt = torch.optim.AdamW([p for p in ps.values() if p.requires_grad], lr=lr, weight_decay=1e-4)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=100, eta_min=1e-5)
    fwd = rp.FORWARDS[kind]
    hist: List[float] = []
    vhist: List[float] = []
    for _ in range(epochs):
        tot = 0.0
        for x, y in batches:
            opt.zero_grad()
            loss = rp.relative_l2(fwd(ps, x, **hyper), y)
            loss.backward()
            opt.step()
            tot += loss.item()
        hist.append(tot / len(batches))
        with torch.no_grad():
            vhist.append(sum(rp.relative_l2(fwd(ps, vx, **hyper), vy).item() for vx, vy in val) / len(val))
        sch.step()
    return hist, vhist, {k: v.detach() for k, v in ps.items()}
