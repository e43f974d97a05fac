"""CPU restatement of the reference's autoregressive rollout and its evaluation loop -- TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and the benchmark's CPU legs may import this module, and only as the checker or
the timed CPU baseline.  The product (resolution_pde_b200/rollout.py) never imports it.

Follows `utils/autoregressive_step.py`:
  perform_rollout       <- perform_rollout_1d (:284-309); the 2-D case is the same loop on (B, H, W) states (the
                           reference has no 2-D rollout; BASELINE config 5 asks for one)
  rollout_batch_loss    <- the inner loop of evaluate_1d_rollout_all_resolution (:141-196): encode the initial condition,
                           roll out, decode, one RelativeL2Loss per step, mean over steps
  evaluate_rollout      <- total_loss / num_batches over the loader (:128-131, 197-199, 238-240)
Normalisers follow SimpleNormalizer (`dataloaders/ks_naive_markov.py:396-406`), passed as (mean, std, eps) tuples.

Pinned by tests/golden/rollout/*.npz, generated from the UNMODIFIED reference perform_rollout_1d, models and
RelativeL2Loss by oracle/make_golden_rollout.py (tests/test_rollout_oracle.py).
"""
from __future__ import annotations

from typing import Callable, Iterable, Optional, Tuple

import torch

Norm = Optional[Tuple[float, float, float]]


def encode(x: torch.Tensor, n: Norm) -> torch.Tensor:
    return x if n is None else (x - n[0]) / (n[1] + n[2])


def decode(x: torch.Tensor, n: Norm) -> torch.Tensor:
    return x if n is None else x * (n[1] + n[2]) + n[0]


def relative_l2_mean(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """RelativeL2Loss(size_average=True, reduction=True) (`utils/loss.py:31-59`)"""
    b = x.shape[0]
    d = torch.norm(x.reshape(b, -1) - y.reshape(b, -1), 2, 1)
    n = torch.norm(y.reshape(b, -1), 2, 1)
    return torch.mean(d / (n + 1e-8))


def perform_rollout(forward: Callable[[torch.Tensor], torch.Tensor], initial_condition: torch.Tensor, rollout_steps: int,
                    x_norm: Norm = None, y_norm: Norm = None) -> torch.Tensor:
    """initial_condition (B, *spatial), already normalised -> normalised predictions (B, T, *spatial)"""
    preds = []
    cur = initial_condition
    for _ in range(rollout_steps):
        nxt = forward(cur.unsqueeze(1))
        if nxt.dim() == cur.dim() + 1 and nxt.shape[1] == 1:
            nxt = nxt.squeeze(1)
        preds.append(nxt.unsqueeze(1))
        if x_norm is not None and y_norm is not None:          # decode then re-encode for the next step (:302-304)
            cur = encode(decode(nxt, y_norm), x_norm)
        else:
            cur = nxt
    return torch.cat(preds, dim=1)


def rollout_batch_loss(forward, trajectories: torch.Tensor, rollout_steps: int, x_norm: Norm = None, y_norm: Norm = None):
    """trajectories (B, T_all, *spatial) un-normalised -> (batch_loss, step_losses (T,), normalised preds)"""
    T = min(rollout_steps, trajectories.shape[1] - 1)
    if T <= 0:
        raise ValueError("not enough timesteps for rollout")
    ic = encode(trajectories[:, 0], x_norm)
    preds = perform_rollout(forward, ic, T, x_norm, y_norm)
    den = decode(preds, y_norm)
    truth = trajectories[:, 1:T + 1]
    steps = torch.stack([relative_l2_mean(den[:, t], truth[:, t]) for t in range(T)])
    return steps.sum() / T, steps, preds


def evaluate_rollout(forward, loader: Iterable[torch.Tensor], rollout_steps: int, x_norm: Norm = None, y_norm: Norm = None) -> float:
    tot, n = 0.0, 0
    for batch in loader:
        if batch.shape[1] - 1 <= 0:
            continue
        tot += float(rollout_batch_loss(forward, batch, rollout_steps, x_norm, y_norm)[0])
        n += 1
    return tot / n
