"""Autoregressive rollout and its evaluation on the device.

The reference rolls a trained one-step model forward in a Python loop (`utils/autoregressive_step.py:284-309`): every
step is `model(state)`, a `torch.cat` bookkeeping entry, `y_normalizer.decode`, `x_normalizer.encode`; the evaluation
around it (`:140-196`) decodes the whole rollout again and then calls `RelativeL2Loss` and `.item()` once per step.
With a fast model these tiny launches and the per-step host syncs are the whole cost.  Here

* the state stays in HBM: one pass per step (`rpde_rollout_renorm`) stores the prediction into the (B, T, ...) trajectory
  and writes the re-normalised next input;
* every step's relative L2 against the ground truth comes from one launch pair over the whole trajectory
  (`rpde_rollout_loss`), reading the ground truth through its strides (no `[:, 1:T+1]` copy) and decoding in registers;
* `RolloutEvaluator` captures the T model forwards, the T re-normalisations and the loss into ONE CUDA graph per
  (trajectory shape, T): a rollout batch is a single replay and no host sync.

`perform_rollout_1d` keeps the reference's name, arguments and return value; `perform_rollout_2d` is the 2-D rollout the
reference lacks and BASELINE config 5 asks for (same semantics, (B, H, W) states).  fp32 CUDA tensors only; the model is a
drop-in module of this package (or anything mapping (B, 1, ...) -> (B, 1, ...) on the device).
"""
from __future__ import annotations

from typing import Dict, Iterable, Optional, Tuple

import torch

from . import _lib
from .ops import _Ctx, _need_cuda, _p


class SimpleNormalizer:
    """Global mean / std normaliser of the reference's dataloaders (`dataloaders/ks_naive_markov.py:396-406`; defined
    inline there, so restated here): encode (x - mean) / (std + eps), decode x (std + eps) + mean, on the device."""

    def __init__(self, mean, std, eps: float = 1e-8):
        self.mean = float(mean)
        self.std = float(std)
        self.eps = eps

    def encode(self, x: torch.Tensor) -> torch.Tensor:
        out = torch.empty_like(x, memory_format=torch.contiguous_format)
        _renorm_launch(x, None, 0, out, 1.0, 0.0, self.mean, self.std + self.eps, True)
        return out

    def decode(self, x: torch.Tensor, device="cuda") -> torch.Tensor:
        out = torch.empty_like(x, memory_format=torch.contiguous_format)
        _renorm_launch(x, None, 0, out, self.std + self.eps, self.mean, 0.0, 1.0, True)
        return out

    def cuda(self):
        return self

    def cpu(self):
        return self


def _affine(x_normalizer, y_normalizer) -> Optional[Tuple[float, float, float, float]]:
    """(y_scale, y_mean, x_mean, x_scale) of encode_x(decode_y(.)), or None when the reference skips it (:302-306)"""
    if x_normalizer is None or y_normalizer is None:
        return None
    for n in (x_normalizer, y_normalizer):
        if not all(hasattr(n, a) for a in ("mean", "std")):
            raise TypeError("rollout: normalisers must carry scalar `mean` and `std` (SimpleNormalizer semantics)")
    ye, xe = getattr(y_normalizer, "eps", 1e-8), getattr(x_normalizer, "eps", 1e-8)
    return float(y_normalizer.std) + ye, float(y_normalizer.mean), float(x_normalizer.mean), float(x_normalizer.std) + xe


def _renorm_launch(pred, preds, t, next_state, ys, ym, xm, xs, renorm, pred_batch_stride=None):
    _need_cuda(pred, preds, next_state)
    B = next_state.shape[0]
    S = next_state.numel() // B
    if pred_batch_stride is None:
        pred = pred.contiguous()
        pred_batch_stride = S
    if preds is None:
        out_ptr, bs = None, S
    else:
        T = preds.shape[1]
        out_ptr, bs = preds.data_ptr() + 4 * t * S, T * S
    with _Ctx(pred) as c:
        _lib.check(c.lib.rpde_rollout_renorm(_p(pred), out_ptr, _p(next_state), B, S, pred_batch_stride, bs, ys, ym, xm,
                                             xs, int(renorm), c.stream), "rpde_rollout_renorm")


def rollout_losses(preds: torch.Tensor, trajectories: torch.Tensor, y_normalizer=None,
                   first_step: int = 1) -> Tuple[torch.Tensor, torch.Tensor]:
    """(batch_loss, step_losses) of a normalised rollout `preds` (B, T, ...) against `trajectories[:, first_step:first_step+T]`
    (B, T_all, ...): step_losses[t] = RelativeL2Loss()(decode(preds[:, t]), truth[:, t]), batch_loss = their mean
    (`autoregressive_step.py:176-196`).  Both stay on the device."""
    _need_cuda(preds, trajectories)
    preds = preds.contiguous()
    trajectories = trajectories.contiguous()
    B, T = preds.shape[0], preds.shape[1]
    S = preds.numel() // (B * T)
    T_all = trajectories.shape[1]
    if trajectories.shape[0] != B or trajectories.numel() != B * T_all * S or first_step + T > T_all:
        raise RuntimeError(f"rollout_losses: trajectories {tuple(trajectories.shape)} do not cover {T} steps of {tuple(preds.shape)}")
    acc = torch.empty(2 * B * T, dtype=torch.float32, device=preds.device)
    steps = torch.empty(T, dtype=torch.float32, device=preds.device)
    total = torch.empty((), dtype=torch.float32, device=preds.device)
    if y_normalizer is None:
        ys, ym, dec = 1.0, 0.0, 0
    else:
        ys, ym, dec = float(y_normalizer.std) + getattr(y_normalizer, "eps", 1e-8), float(y_normalizer.mean), 1
    with _Ctx(preds) as c:
        _lib.check(c.lib.rpde_rollout_loss(_p(preds), trajectories.data_ptr() + 4 * first_step * S, _p(acc), _p(steps),
                                           _p(total), B, T, S, T * S, S, T_all * S, S, ys, ym, dec, c.stream),
                   "rpde_rollout_loss")
    return total, steps


def _rollout(model, state: torch.Tensor, rollout_steps: int, x_normalizer, y_normalizer) -> torch.Tensor:
    """state (B, *spatial) normalised -> normalised predictions (B, T, *spatial)"""
    _need_cuda(state)
    if rollout_steps <= 0:
        raise ValueError("rollout_steps must be positive")
    aff = _affine(x_normalizer, y_normalizer)
    ys, ym, xm, xs = aff if aff is not None else (1.0, 0.0, 0.0, 1.0)
    B, spatial = state.shape[0], tuple(state.shape[1:])
    preds = torch.empty((B, rollout_steps) + spatial, dtype=torch.float32, device=state.device)
    cur = state.contiguous()
    nxt = [torch.empty_like(cur), torch.empty_like(cur)]
    with torch.no_grad():
        for t in range(rollout_steps):
            out = model(cur.unsqueeze(1))                      # (B, 1, *spatial), as the reference feeds it (:290-295)
            if out.shape[0] != B or out.numel() != cur.numel():
                raise RuntimeError(f"rollout: the model maps {tuple(cur.unsqueeze(1).shape)} to {tuple(out.shape)}; "
                                   "a one-channel state-to-state model is required")
            _renorm_launch(out, preds, t, nxt[t & 1], ys, ym, xm, xs, aff is not None)
            cur = nxt[t & 1]
    return preds


def perform_rollout_1d(model, initial_condition, rollout_steps, model_type="ffno1d", time_val=None, device="cuda",
                       x_normalizer=None, y_normalizer=None):
    """Drop-in for `utils/autoregressive_step.py:284-309`: initial_condition (B, N) already normalised ->
    normalised predictions (B, rollout_steps, N)."""
    if model_type == "pos":
        raise NotImplementedError("perform_rollout_1d: the 'pos' (time-conditioned) model is outside the spectral path")
    if initial_condition.dim() != 2:
        raise RuntimeError(f"perform_rollout_1d expects (batch, spatial), got {tuple(initial_condition.shape)}")
    return _rollout(model, initial_condition, int(rollout_steps), x_normalizer, y_normalizer)


def perform_rollout_2d(model, initial_condition, rollout_steps, model_type="ffno2d", time_val=None, device="cuda",
                       x_normalizer=None, y_normalizer=None):
    """2-D counterpart (BASELINE config 5; the reference has none): (B, H, W) -> (B, rollout_steps, H, W)."""
    if initial_condition.dim() != 3:
        raise RuntimeError(f"perform_rollout_2d expects (batch, H, W), got {tuple(initial_condition.shape)}")
    return _rollout(model, initial_condition, int(rollout_steps), x_normalizer, y_normalizer)


class RolloutEvaluator:
    """evaluator(trajectories) -> (batch_loss, step_losses, preds): one rollout batch of `evaluate_1d_rollout_all_resolution`'s
    inner loop (`autoregressive_step.py:141-196`) as one CUDA-graph replay.

    trajectories: (B, T_all, N) or (B, T_all, H, W), un-normalised, on the device or in pinned host memory.  The initial
    condition trajectories[:, 0] is encoded with x_normalizer, rolled out min(rollout_steps, T_all - 1) steps, decoded with
    y_normalizer and compared with trajectories[:, 1:T+1].  The returned tensors are static buffers of the graph (valid
    until the next call with the same shape); `preds` holds the NORMALISED predictions like perform_rollout_1d.
    """

    def __init__(self, model: torch.nn.Module, rollout_steps: int, x_normalizer=None, y_normalizer=None, warmup: int = 1,
                 use_graph: bool = True):
        self.model = model
        self.rollout_steps = int(rollout_steps)
        self.x_normalizer = x_normalizer
        self.y_normalizer = y_normalizer
        self.warmup = warmup
        self.use_graph = use_graph
        self._graphs: Dict[Tuple, Tuple] = {}

    def _run(self, traj: torch.Tensor, ic: torch.Tensor, T: int):
        xn = self.x_normalizer
        xm, xs = (float(xn.mean), float(xn.std) + getattr(xn, "eps", 1e-8)) if xn is not None else (0.0, 1.0)
        # trajectories[:, 0] read through its batch stride: encode (or plain gather) in one pass, no torch copy
        _renorm_launch(traj, None, 0, ic, 1.0, 0.0, xm, xs, xn is not None, pred_batch_stride=traj[0].numel())
        preds = _rollout(self.model, ic, T, self.x_normalizer, self.y_normalizer)
        total, steps = rollout_losses(preds, traj, self.y_normalizer, 1)
        return total, steps, preds

    def _capture(self, shape):
        dev = next(self.model.parameters()).device
        T = min(self.rollout_steps, shape[1] - 1)
        if T <= 0:
            raise ValueError(f"not enough timesteps ({shape[1]}) for rollout")        # the reference skips the batch (:149-151)
        traj = torch.zeros(shape, dtype=torch.float32, device=dev)
        ic = torch.empty((shape[0],) + tuple(shape[2:]), dtype=torch.float32, device=dev)
        if not self.use_graph:
            return None, traj, ic, T, None
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                                                 # builds tables / grids off the capture
            for _ in range(max(1, self.warmup)):
                self._run(traj, ic, T)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = self._run(traj, ic, T)
        return g, traj, ic, T, out

    def __call__(self, trajectories: torch.Tensor):
        if trajectories.dim() not in (3, 4):
            raise RuntimeError(f"RolloutEvaluator expects (B, T, N) or (B, T, H, W), got {tuple(trajectories.shape)}")
        key = tuple(trajectories.shape)
        ent = self._graphs.get(key)
        if ent is None:
            ent = self._capture(key)
            self._graphs[key] = ent
        g, traj, ic, T, out = ent
        if trajectories.data_ptr() != traj.data_ptr():
            traj.copy_(trajectories, non_blocking=True)
        if g is None:
            return self._run(traj, ic, T)
        g.replay()
        return out


def mean_over_batches(trajectory_loader: Iterable[torch.Tensor], batch_loss, device, distributed: bool = False, group=None) -> float:
    """total_loss / num_batches of `autoregressive_step.py:128-131, 197-199, 238-240` with the batch losses kept on the
    device: one host sync at the end instead of one per step.  Batches too short to roll out are skipped as there.
    distributed=True ("rollout/eval sweeps shard trajectories/batches across ranks"): rank r takes the valid
    batches r, r + world, ...; one all-reduce of (sum of batch losses, batch count) at the end, so every rank returns the
    single-process mean.  No collective inside the rollout itself."""
    import torch.distributed as dist
    on = distributed and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    rank, world = (dist.get_rank(group), dist.get_world_size(group)) if on else (0, 1)
    acc = torch.zeros(2, dtype=torch.float32, device=device)      # (sum of batch losses, number of batches)
    i = 0
    for batch in trajectory_loader:
        if batch.shape[1] - 1 <= 0:
            continue
        if i % world == rank:
            acc[0] += batch_loss(batch)
            acc[1] += 1
        i += 1
    if on:
        dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
    tot, n = acc.tolist()
    if n == 0:
        raise ValueError("evaluate_rollout: no valid batches")
    return tot / n


def evaluate_rollout(model, trajectory_loader: Iterable[torch.Tensor], rollout_steps: int = 10, x_normalizer=None,
                     y_normalizer=None, use_graph: bool = True, distributed: bool = False, group=None,
                     evaluator: Optional[RolloutEvaluator] = None) -> float:
    """Mean over batches of the per-batch rollout loss, the number `evaluate_1d_rollout_all_resolution` stores per
    resolution; every batch is one CUDA-graph replay (`RolloutEvaluator`; pass `evaluator` to reuse its captured graphs across
    calls), see `mean_over_batches` for the multi-GPU split."""
    ev = evaluator or RolloutEvaluator(model, rollout_steps, x_normalizer, y_normalizer, use_graph=use_graph)
    dev = next(model.parameters()).device
    return mean_over_batches(trajectory_loader, lambda b: ev(b.to(torch.float32))[0], dev, distributed, group)
