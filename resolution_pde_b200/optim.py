"""Fused AdamW for the spectral models.

The reference builds `torch.optim.AdamW(model.parameters(), lr=..., weight_decay=...)` (main_1d.py:144,
main_2d.py:173) and steps it eagerly: one foreach launch group per state tensor list, bias corrections on the host.
`FusedAdamW` keeps the constructor, the `param_groups` / `state_dict` layout (`step`, `exp_avg`, `exp_avg_sq`) and the
arithmetic, but updates every parameter of a group in ONE kernel (`rpde_adamw_step`): a device table of
(param, grad, exp_avg, exp_avg_sq, numel) records, lr and the step count as device scalars.  Nothing in `step()`
synchronises or depends on host values, so it can be captured inside `GraphedTrainStep`; a scheduler changing
`group["lr"]` is picked up by a 4-byte device copy before the next replay.

Complex parameters (the FNO spectral weights) are updated through their (re, im) float view, which is what
torch.optim.AdamW does (`torch.view_as_real`).  amsgrad / maximize / differentiable are not supported (the reference
does not use them) and raise at construction.
"""
from __future__ import annotations

import math
from typing import Iterable

import torch

from . import _lib


def _real(t: torch.Tensor) -> torch.Tensor:
    return torch.view_as_real(t) if t.is_complex() else t


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params: Iterable, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-2,
                 amsgrad: bool = False, *, maximize: bool = False):
        if amsgrad or maximize:
            raise ValueError("FusedAdamW: amsgrad / maximize are not supported")
        if lr < 0 or eps < 0 or weight_decay < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1):
            raise ValueError("FusedAdamW: invalid hyper-parameters")
        # the extra keys make state_dict()["param_groups"] loadable by torch.optim.AdamW (and the other way round)
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, amsgrad=False,
                                      maximize=False, foreach=None, capturable=False, differentiable=False, fused=None,
                                      decoupled_weight_decay=True))
        self._groups = {}   # id(group) -> cached device table

    # ------------------------------------------------------------------------------------------
    def _state_for(self, p: torch.Tensor):
        st = self.state[p]
        if not st:
            st["step"] = torch.zeros((), dtype=torch.float32)            # host copy, torch's state_dict layout
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    def _table(self, group):
        """Device table of a group for the current (param, grad) pointers.  Tables are cached per pointer set and never
        freed: a captured graph keeps reading the table it was captured with (one per batch shape in GraphedTrainStep)."""
        ps = [p for p in group["params"] if p.grad is not None]
        if not ps:
            return None, None
        ent = self._groups.get(id(group))
        key = tuple((p.data_ptr(), p.grad.data_ptr()) for p in ps)
        if ent is not None and key in ent["tables"]:
            return ent, ent["tables"][key]
        for p in ps:
            if not p.is_cuda or _real(p).dtype != torch.float32:
                raise RuntimeError("FusedAdamW: parameters must be fp32 / complex64 CUDA tensors (no CPU fallback)")
            if not p.is_contiguous() or not p.grad.is_contiguous():
                raise RuntimeError("FusedAdamW: parameters and gradients must be contiguous")
        dev = ps[0].device
        rows, chunk0 = [], 0
        for p in ps:
            st = self._state_for(p)
            n = _real(p).numel()
            rows.append([p.data_ptr(), p.grad.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), n, chunk0])
            chunk0 += (n + 1023) // 1024
        if ent is None:
            step0 = max((float(self.state[p]["step"]) for p in ps), default=0.0)
            ent = dict(hyper=torch.tensor([group["lr"], step0, 1.0, 1.0], dtype=torch.float32).to(dev), lr=group["lr"], tables={},
                       params=ps)
            self._groups[id(group)] = ent
        tab = dict(table=torch.tensor(rows, dtype=torch.int64).to(dev), n=len(ps), chunks=chunk0)
        ent["tables"][key] = tab
        return ent, tab

    def sync_lr(self):
        """Push a changed group["lr"] (scheduler) to the device scalar.  Outside any capture."""
        for group in self.param_groups:
            ent = self._groups.get(id(group))
            if ent is not None and ent["lr"] != group["lr"]:
                ent["hyper"][0:1].fill_(group["lr"])
                ent["lr"] = group["lr"]

    def launch(self):
        """Enqueue the update kernels on the current stream (no host-dependent values: graph-capturable once the
        tables for the current gradient tensors exist, i.e. after one eager `step()` / `prepare()`)."""
        lib = _lib.load()
        for group in self.param_groups:
            ent, tab = self._table(group)
            if ent is None:
                continue
            b1, b2 = group["betas"]
            dev = tab["table"].device
            with torch.cuda.device(dev):
                rc = lib.rpde_adamw_step(tab["table"].data_ptr(), tab["n"], tab["chunks"], ent["hyper"].data_ptr(),
                                         b1, b2, group["eps"], group["weight_decay"],
                                         torch.cuda.current_stream(dev).cuda_stream)
            _lib.check(rc, "rpde_adamw_step")

    def prepare(self):
        """Build the device tables for the gradients currently attached to the parameters (host -> device copies)."""
        for group in self.param_groups:
            self._table(group)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self.prepare()
        self.sync_lr()
        self.launch()
        return loss

    def state_dict(self):
        # bring the device step counts back so the dict matches torch.optim.AdamW's (one sync, checkpoint time only)
        for group in self.param_groups:
            ent = self._groups.get(id(group))
            if ent is not None:
                t = float(ent["hyper"][1].item())
                for p in ent["params"]:
                    self.state[p]["step"] = torch.tensor(t, dtype=torch.float32)
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._groups.clear()                                    # tables and device step counts are rebuilt lazily
