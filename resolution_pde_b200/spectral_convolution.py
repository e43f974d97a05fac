"""Spectral convolution layers (reference: models/spectral_convolution.py:24-98, 108-318).

Same class names, constructor arguments, forward signatures / return types and parameter layout
(state-dict compatible: complex64 `weights1/weights2`, float32 `(C, C, modes, 2)` `fourier_weight.N`).
The arithmetic -- truncated forward DFT, per-mode complex channel mixing, truncated inverse DFT, and
their adjoints -- runs in librpde_b200.so; torch.fft and torch.einsum are never called.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .custom_layer import FeedForward


class SpectralConv1d(nn.Module):
    """spectral_convolution.py:24-55: rfft -> first `modes1` bins times complex weights -> irfft.
    (B, in, N) -> (B, out, N) at any N with modes1 <= N//2+1 (no clamp: more modes raises RuntimeError)."""

    def __init__(self, in_channels, out_channels, modes1):
        super().__init__()
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.modes1 = modes1
        self.scale = (1 / (in_channels * out_channels))
        self.weights1 = nn.Parameter(self.scale * torch.rand(in_channels, out_channels, self.modes1, dtype=torch.cfloat))

    def forward(self, x):
        return ops.fno_spectral(x, self.weights1)


class SpectralConv2d(nn.Module):
    """spectral_convolution.py:61-98: rfft2 -> rows [:m1] x cols [:m2] with weights1, rows [-m1:] with
    weights2 -> irfft2.  (B, in, H, W) -> (B, out, H, W), non-square and odd sizes included."""

    def __init__(self, in_channels, out_channels, modes1, modes2):
        super().__init__()
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.modes1 = modes1
        self.modes2 = modes2
        self.scale = (1 / (in_channels * out_channels))
        self.weights1 = nn.Parameter(
            self.scale * torch.rand(in_channels, out_channels, self.modes1, self.modes2, dtype=torch.cfloat))
        self.weights2 = nn.Parameter(
            self.scale * torch.rand(in_channels, out_channels, self.modes1, self.modes2, dtype=torch.cfloat))

    def forward(self, x):
        return ops.fno_spectral(x, self.weights1, self.weights2)


act_registry = {"gelu": F.gelu, "identity": nn.Identity(), "relu": F.relu}
_act_ids = {"gelu": ops.ACT_GELU, "identity": ops.ACT_NONE, "relu": ops.ACT_RELU}


def _init_fourier_weights(d_model, modes, n):
    ws = nn.ParameterList([])
    for _ in range(n):
        weight = torch.empty(d_model, d_model, modes, 2, dtype=torch.float32)
        param = nn.Parameter(weight)
        nn.init.xavier_normal_(param)
        ws.append(param)
    return ws


def _route_skip(x):
    """True when the layer's skip connection should travel through the Fourier function's second output, so that its gradient
    is accumulated inside the spectral backward pass instead of by a separate add (only matters when a backward pass follows)."""
    return torch.is_grad_enabled() and x.requires_grad and x.is_cuda


class FSpectralConv1d(nn.Module):
    """spectral_convolution.py:108-204.  forward(x (B,N,C), batch_dt=None) -> (act(backcast_ff(fourier(x))), None)."""

    def __init__(self, d_model, modes, forecast_ff=None, backcast_ff=None, fourier_weight=None, factor=4,
                 ff_weight_norm=False, n_ff_layers=2, layer_norm=False, use_fork=False, dropout=0.0, mode='full',
                 activation="identity", fft_norm="ortho", **kwargs):
        super().__init__()
        self.in_dim = d_model
        self.out_dim = d_model
        self.n_modes = modes
        self.mode = mode
        self.use_fork = use_fork
        self.fft_norm = fft_norm
        self.fourier_weight = fourier_weight
        if not self.fourier_weight:
            self.fourier_weight = _init_fourier_weights(d_model, modes, 1)
        self.backcast_ff = backcast_ff
        if not self.backcast_ff:
            self.backcast_ff = FeedForward(d_model, factor, ff_weight_norm=ff_weight_norm, n_layers=n_ff_layers,
                                           layer_norm=layer_norm, dropout=dropout)
        self.act = act_registry[activation]
        self._act_id = _act_ids[activation]

    def forward(self, x, batch_dt=None, residual=None):
        """`residual` (not in the reference signature, default None) lets FFNO1D fuse its skip connection."""
        if self.mode != 'no-fourier':
            if residual is x and _route_skip(x):
                x, residual = self.forward_fourier(x, with_skip=True)
            else:
                x = self.forward_fourier(x)
        if isinstance(self.backcast_ff, FeedForward):
            return self.backcast_ff(x, residual=residual, final_act=self._act_id), None
        b = self.act(self.backcast_ff(x))  # injected foreign feed-forward module
        return (b if residual is None else residual + b), None

    def forward_fourier(self, x, with_skip=False):
        if self.mode == 'full':
            mix = ops.MIX_FULL
        elif self.mode == 'low-pass':
            mix = ops.MIX_LOWPASS
        else:
            raise ValueError(f"Mode {self.mode} not recognized")
        return ops.ffno_fourier(x, [self.fourier_weight[0]], self.n_modes, mix, self.fft_norm, with_skip=with_skip)


class FSpectralConv2d(nn.Module):
    """spectral_convolution.py:210-318.  forward(x (B,M,N,C), batch_dt=None) -> (backcast_ff(.), forecast or None).
    fourier_weight[0] acts along the last spatial axis, [1] along axis -2; both passes read the same input and
    their results are summed; 'ortho' normalisation is hard-coded as in the reference."""

    def __init__(self, d_model, modes, forecast_ff=None, backcast_ff=None, fourier_weight=None, factor=4,
                 ff_weight_norm=False, n_ff_layers=2, layer_norm=False, use_fork=False, dropout=0.0, mode='full'):
        super().__init__()
        self.in_dim = d_model
        self.out_dim = d_model
        self.n_modes = modes
        self.mode = mode
        self.use_fork = use_fork
        self.fourier_weight = fourier_weight
        if not self.fourier_weight:
            self.fourier_weight = _init_fourier_weights(d_model, modes, 2)
        self.backcast_ff = backcast_ff
        if not self.backcast_ff:
            self.backcast_ff = FeedForward(d_model, factor, ff_weight_norm=ff_weight_norm, n_layers=n_ff_layers,
                                           layer_norm=layer_norm, dropout=dropout)
        self.forecast_ff = forecast_ff
        if use_fork and not self.forecast_ff:
            self.forecast_ff = FeedForward(d_model, factor, ff_weight_norm=ff_weight_norm, n_layers=n_ff_layers,
                                           layer_norm=layer_norm, dropout=dropout)

    def forward(self, x, batch_dt=None, residual=None):
        if self.mode != 'no-fourier':
            if residual is x and self.mode in ('full', 'low-pass') and _route_skip(x):
                x, residual = self.forward_fourier(x, with_skip=True)
            else:
                x = self.forward_fourier(x)
        if isinstance(self.backcast_ff, FeedForward):
            b = self.backcast_ff(x, residual=residual)
        else:
            b = self.backcast_ff(x)
            b = b if residual is None else residual + b
        f = self.forecast_ff(x) if self.use_fork else None
        return b, f

    def forward_fourier(self, x, with_skip=False):
        if self.mode == 'full':
            mix = ops.MIX_FULL
        elif self.mode == 'low-pass':
            mix = ops.MIX_LOWPASS
        else:
            # the reference has no `else` branch: both spectra stay zero (spectral_convolution.py:273-281)
            return torch.zeros_like(x)
        return ops.ffno_fourier(x, [self.fourier_weight[0], self.fourier_weight[1]], self.n_modes, mix, "ortho",
                                with_skip=with_skip)
