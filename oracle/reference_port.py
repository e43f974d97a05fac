"""CPU oracle for the spectral-convolution hot path  --  TEST INFRASTRUCTURE ONLY.

This file is a from-scratch *functional* restatement (torch on the CPU, any
float dtype, fp64 for arbitration) of the algorithm the reference implements in

    models/spectral_convolution.py:24-98, 108-318   (the four spectral layers)
    models/fno_blocks.py:25-83, models/fno.py:24-150 (FNO1d / FNO2d)
    models/custom_layer.py:49-108, models/ffno.py:25-237 (FeedForward, WNLinear, FFNO1D/2D)
    utils/loss.py:17-59                              (RelativeL2Loss)

It takes the reference's *state-dict* (same key names, shapes, dtypes) plus the
constructor hyper-parameters and evaluates the model as plain functions, so
autograd on the CPU gives oracle gradients for every parameter.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module, and only as the checker or the timed CPU
baseline.  The product path (resolution_pde_b200/) never imports it.

Parity pin: the reference has no tests or golden vectors of its own (SURVEY §4,
§8c).  The pin is therefore tests/golden/*.npz, produced by oracle/make_golden.py
from the UNMODIFIED reference modules imported from /root/reference in the build
container; tests/test_oracle_golden.py checks this port against those vectors.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence

import torch
import torch.nn.functional as F

Params = Dict[str, torch.Tensor]


# --------------------------------------------------------------------------
# small helpers
# --------------------------------------------------------------------------
def _act(name_or_fn):
    """activation registry of the reference (spectral_convolution.py:104-106)."""
    if callable(name_or_fn):
        return name_or_fn
    table = {"gelu": F.gelu, "identity": (lambda t: t), "relu": F.relu}
    return table[name_or_fn]


def _sub(params: Params, prefix: str) -> Params:
    n = len(prefix)
    return {k[n:]: v for k, v in params.items() if k.startswith(prefix)}


def _wn_weight(p: Params) -> torch.Tensor:
    """WNLinear (custom_layer.py:70-108): legacy weight_norm, dim=0:
    w = g * v / ||v||_row.  Plain Linear when no weight_g/weight_v present."""
    if "weight" in p:
        return p["weight"]
    v, g = p["weight_v"], p["weight_g"]
    return v * (g / v.norm(dim=1, keepdim=True))


# --------------------------------------------------------------------------
# FNO spectral layers (channels-first, complex weights, norm="backward")
# --------------------------------------------------------------------------
def spectral_conv1d(x: torch.Tensor, weights1: torch.Tensor) -> torch.Tensor:
    """spectral_convolution.py:38-55.  x (B,I,N); weights1 complex (I,O,modes).
    No clamp of modes: modes > N//2+1 is an error, as in the reference."""
    B, I, N = x.shape
    modes = weights1.shape[2]
    if modes > N // 2 + 1:
        raise RuntimeError("SpectralConv1d: modes1 exceeds N//2+1 (reference einsum shape error)")
    xf = torch.fft.rfft(x)
    out = torch.zeros(B, weights1.shape[1], N // 2 + 1, dtype=xf.dtype)
    out[:, :, :modes] = torch.einsum("bik,iok->bok", xf[:, :, :modes], weights1.to(xf.dtype))
    return torch.fft.irfft(out, n=N)


def spectral_conv2d(x: torch.Tensor, weights1: torch.Tensor, weights2: torch.Tensor) -> torch.Tensor:
    """spectral_convolution.py:79-98.  x (B,I,H,W); weights complex (I,O,m1,m2).
    Rows [:m1] use weights1, rows [-m1:] use weights2, columns [:m2]."""
    B, I, H, W = x.shape
    m1, m2 = weights1.shape[2], weights1.shape[3]
    if m2 > W // 2 + 1 or m1 > H:
        raise RuntimeError("SpectralConv2d: modes exceed the available spectrum")
    xf = torch.fft.rfft2(x)
    out = torch.zeros(B, weights1.shape[1], H, W // 2 + 1, dtype=xf.dtype)
    out[:, :, :m1, :m2] = torch.einsum("bixy,ioxy->boxy", xf[:, :, :m1, :m2], weights1.to(xf.dtype))
    out[:, :, H - m1:, :m2] = torch.einsum("bixy,ioxy->boxy", xf[:, :, H - m1:, :m2], weights2.to(xf.dtype))
    return torch.fft.irfft2(out, s=(H, W))


# --------------------------------------------------------------------------
# F-FNO factorised spectral layers (channels-last, real (C,C,modes,2) weights)
# --------------------------------------------------------------------------
def fspectral_fourier1d(x: torch.Tensor, w: torch.Tensor, n_modes: int, mode: str = "full",
                        fft_norm: str = "ortho") -> torch.Tensor:
    """spectral_convolution.py:159-204.  x (B,N,C) channels-last."""
    B, N, C = x.shape
    xt = x.transpose(1, 2)
    xf = torch.fft.rfft(xt, dim=-1, norm=fft_norm)
    K = min(n_modes, N // 2 + 1)
    out = torch.zeros_like(xf)
    if mode == "full":
        wc = torch.view_as_complex(w[:, :, :K].contiguous())
        out[:, :, :K] = torch.einsum("bik,iok->bok", xf[:, :, :K], wc.to(xf.dtype))
    elif mode == "low-pass":
        out[:, :, :K] = xf[:, :, :K]
    else:
        raise ValueError(f"Mode {mode} not recognized")
    y = torch.fft.irfft(out, n=N, dim=-1, norm=fft_norm)
    return y.transpose(1, 2)


def fspectral_fourier2d(x: torch.Tensor, w_y: torch.Tensor, w_x: torch.Tensor, n_modes: int,
                        mode: str = "full") -> torch.Tensor:
    """spectral_convolution.py:256-318.  x (B,M,N,C) channels-last; w_y = fourier_weight[0]
    acts along the LAST spatial axis (N), w_x = fourier_weight[1] along axis -2 (M); 'ortho'
    hard-coded; both passes read the same input; unknown mode silently gives zeros."""
    B, M, N, C = x.shape
    xc = x.permute(0, 3, 1, 2)
    # y pass
    fy = torch.fft.rfft(xc, dim=-1, norm="ortho")
    Ky = min(n_modes, N // 2 + 1)
    oy = torch.zeros_like(fy)
    if mode == "full":
        wc = torch.view_as_complex(w_y[:, :, :Ky].contiguous())
        oy[..., :Ky] = torch.einsum("bixy,ioy->boxy", fy[..., :Ky], wc.to(fy.dtype))
    elif mode == "low-pass":
        oy[..., :Ky] = fy[..., :Ky]
    xy = torch.fft.irfft(oy, n=N, dim=-1, norm="ortho")
    # x pass
    fx = torch.fft.rfft(xc, dim=-2, norm="ortho")
    Kx = min(n_modes, M // 2 + 1)
    ox = torch.zeros_like(fx)
    if mode == "full":
        wc = torch.view_as_complex(w_x[:, :, :Kx].contiguous())
        ox[:, :, :Kx, :] = torch.einsum("bixy,iox->boxy", fx[:, :, :Kx, :], wc.to(fx.dtype))
    elif mode == "low-pass":
        ox[:, :, :Kx, :] = fx[:, :, :Kx, :]
    xx = torch.fft.irfft(ox, n=M, dim=-2, norm="ortho")
    return (xx + xy).permute(0, 2, 3, 1)


def feed_forward(x: torch.Tensor, p: Params, n_layers: int, layer_norm: bool) -> torch.Tensor:
    """custom_layer.py:49-68 at dropout p=0: Linear -> GELU (all but last) -> LayerNorm (last only)."""
    for i in range(n_layers):
        x = F.linear(x, p[f"layers.{i}.0.weight"], p[f"layers.{i}.0.bias"])
        if i < n_layers - 1:
            x = F.gelu(x)
        elif layer_norm:
            x = F.layer_norm(x, (x.shape[-1],), p[f"layers.{i}.3.weight"], p[f"layers.{i}.3.bias"])
    return x


def fspectral_conv1d(x, p: Params, n_modes: int, n_ff_layers: int, layer_norm: bool, mode="full",
                     fft_norm="ortho", activation="identity"):
    """FSpectralConv1d.forward (spectral_convolution.py:145-157) -> (tensor, None)."""
    if mode != "no-fourier":
        x = fspectral_fourier1d(x, p["fourier_weight.0"], n_modes, mode, fft_norm)
    b = feed_forward(x, _sub(p, "backcast_ff."), n_ff_layers, layer_norm)
    return _act(activation)(b), None


def fspectral_conv2d(x, p: Params, n_modes: int, n_ff_layers: int, layer_norm: bool, mode="full"):
    """FSpectralConv2d.forward (spectral_convolution.py:248-254), use_fork=False."""
    if mode != "no-fourier":
        x = fspectral_fourier2d(x, p["fourier_weight.0"], p["fourier_weight.1"], n_modes, mode)
    return feed_forward(x, _sub(p, "backcast_ff."), n_ff_layers, layer_norm), None


# --------------------------------------------------------------------------
# whole models (state-dict driven)
# --------------------------------------------------------------------------
def _linspace(a: float, b: float, n: int, dtype) -> torch.Tensor:
    """The reference builds grids with np.linspace in fp64, then casts to fp32
    (fno.py:51, 121-125; ffno.py:92, 201-205)."""
    return torch.linspace(a, b, n, dtype=torch.float64).to(torch.float32).to(dtype)


def fno1d_forward(p: Params, x: torch.Tensor, activation=F.relu, n_blocks: int = 4,
                  grid: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fno.py:56-76 + fno_blocks.py:25-45.  x (B,in,N) -> (B,out,N)."""
    B, _, N = x.shape
    g = _linspace(0.0, 2.0 * math.pi, N, x.dtype) if grid is None else torch.as_tensor(grid, dtype=torch.float32).to(x.dtype)
    x = torch.cat((x, g.reshape(1, 1, N).expand(B, 1, N)), dim=1)
    x = F.conv1d(x, p["lifting.weight"], p["lifting.bias"])
    act = _act(activation)
    for i in range(n_blocks):
        s = spectral_conv1d(x, p[f"fno_blocks.{i}.spectral_conv.weights1"])
        x = act(s + F.conv1d(x, p[f"fno_blocks.{i}.bypass_conv.weight"], p[f"fno_blocks.{i}.bypass_conv.bias"]))
    x = F.gelu(F.conv1d(x, p["projection.mlp1.weight"], p["projection.mlp1.bias"]))
    return F.conv1d(x, p["projection.mlp2.weight"], p["projection.mlp2.bias"])


def fno2d_forward(p: Params, x: torch.Tensor, activation=F.gelu, n_blocks: int = 4,
                  grid: Optional[Sequence[torch.Tensor]] = None) -> torch.Tensor:
    """fno.py:131-150 + fno_blocks.py:63-83.  x (B,in,H,W) -> (B,out,H,W)."""
    B, _, H, W = x.shape
    if grid is None:
        gx, gy = _linspace(0.0, 1.0, H, x.dtype), _linspace(0.0, 1.0, W, x.dtype)
    else:
        gx = torch.as_tensor(grid[0], dtype=torch.float32).to(x.dtype)
        gy = torch.as_tensor(grid[1], dtype=torch.float32).to(x.dtype)
    x = torch.cat((x, gx.reshape(1, 1, H, 1).expand(B, 1, H, W), gy.reshape(1, 1, 1, W).expand(B, 1, H, W)), dim=1)
    x = F.conv2d(x, p["lifting.weight"], p["lifting.bias"])
    act = _act(activation)
    for i in range(n_blocks):
        s = spectral_conv2d(x, p[f"fno_blocks.{i}.spectral_conv.weights1"], p[f"fno_blocks.{i}.spectral_conv.weights2"])
        x = act(s + F.conv2d(x, p[f"fno_blocks.{i}.bypass_conv.weight"], p[f"fno_blocks.{i}.bypass_conv.bias"]))
    x = F.gelu(F.conv2d(x, p["projection.mlp1.weight"], p["projection.mlp1.bias"]))
    return F.conv2d(x, p["projection.mlp2.weight"], p["projection.mlp2.bias"])


def ffno1d_forward(p: Params, x: torch.Tensor, n_layers=4, n_modes=16, n_ff_layers=2, layer_norm=False,
                   grid=None, mode="full", fft_norm="ortho", activation="identity") -> torch.Tensor:
    """ffno.py:98-125.  Quirk kept (ffno.py:51): use_grid := grid, so with grid=None no grid
    channel is concatenated whatever `use_grid` says."""
    B, _, N = x.shape
    if grid is not None:
        g = torch.as_tensor(grid, dtype=torch.float32).to(x.dtype)
        x = torch.cat((x, g.reshape(1, 1, N).expand(B, 1, N)), dim=1)
    x = x.permute(0, 2, 1)
    ip = _sub(p, "in_proj.")
    x = F.linear(x, _wn_weight(ip), ip["bias"])
    for i in range(n_layers):
        b, _ = fspectral_conv1d(x, _sub(p, f"fourier_layers.{i}."), n_modes, n_ff_layers, layer_norm, mode,
                                fft_norm, activation)
        x = x + b
    op = _sub(p, "out_proj.")
    x = F.linear(x, _wn_weight(op), op["bias"])
    return x.permute(0, 2, 1)


def ffno2d_forward(p: Params, x: torch.Tensor, n_layers=4, n_modes=16, n_ff_layers=2, layer_norm=False,
                   grid=None, mode="full", use_grid=True) -> torch.Tensor:
    """ffno.py:210-237.  x (B,in,M,N) -> (B,out,M,N); grid linspace(0,1) per axis."""
    B, _, M, N = x.shape
    if use_grid:
        if grid is None:
            gx, gy = _linspace(0.0, 1.0, M, x.dtype), _linspace(0.0, 1.0, N, x.dtype)
        else:
            gx = torch.as_tensor(grid[0], dtype=torch.float32).to(x.dtype)
            gy = torch.as_tensor(grid[1], dtype=torch.float32).to(x.dtype)
        x = torch.cat((x, gx.reshape(1, 1, M, 1).expand(B, 1, M, N), gy.reshape(1, 1, 1, N).expand(B, 1, M, N)), dim=1)
    x = x.permute(0, 2, 3, 1)
    ip = _sub(p, "in_proj.")
    x = F.linear(x, _wn_weight(ip), ip["bias"])
    for i in range(n_layers):
        b, _ = fspectral_conv2d(x, _sub(p, f"fourier_layers.{i}."), n_modes, n_ff_layers, layer_norm, mode)
        x = x + b
    op = _sub(p, "out_proj.")
    x = F.linear(x, _wn_weight(op), op["bias"])
    return x.permute(0, 3, 1, 2)


def relative_l2(x: torch.Tensor, y: torch.Tensor, size_average: bool = True, reduction: bool = True) -> torch.Tensor:
    """utils/loss.py:31-59: per-sample ||x-y||_2 / (||y||_2 + 1e-8), then mean / sum / none."""
    n = x.shape[0]
    d = (x.reshape(n, -1) - y.reshape(n, -1)).norm(2, 1)
    r = d / (y.reshape(n, -1).norm(2, 1) + 1e-8)
    if not reduction:
        return r
    return r.mean() if size_average else r.sum()


FORWARDS = {"FNO1d": fno1d_forward, "FNO2d": fno2d_forward, "FFNO1D": ffno1d_forward, "FFNO2D": ffno2d_forward}


def step_loss_and_grads(kind: str, params: Params, x: torch.Tensor, y: torch.Tensor, dtype=torch.float32, **hyper):
    """One fwd + RelativeL2 + bwd on the CPU.  Returns (pred, loss, {name: grad}) with the
    parameters cast to `dtype` (fp64 = arbiter)."""
    def cast(t):
        if t.is_complex():
            return t.to(torch.complex128 if dtype == torch.float64 else torch.complex64)
        return t.to(dtype)
    ps = {k: cast(v.detach()).requires_grad_(True) for k, v in params.items()}
    pred = FORWARDS[kind](ps, x.to(dtype), **hyper)
    loss = relative_l2(pred, y.to(dtype))
    names = list(ps)
    grads = torch.autograd.grad(loss, [ps[k] for k in names], allow_unused=True)
    return pred.detach(), loss.detach(), {k: (g if g is not None else torch.zeros_like(ps[k])) for k, g in zip(names, grads)}
