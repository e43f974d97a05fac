"""FFNO1D / FFNO2D (reference: models/ffno.py:25-237): same constructors, forward signatures, state-dict."""
from __future__ import annotations

import torch.nn as nn

from . import ops
from .custom_layer import Linear, WNLinear
from .spectral_convolution import FSpectralConv1d, FSpectralConv2d


class FFNO1D(nn.Module):
    def __init__(self, in_channels, out_channels, width=64, n_layers=4, n_modes=16, factor=4, ff_weight_norm=False,
                 n_ff_layers=2, layer_norm=False, dropout=0.0, grid=None, mode='full', fft_norm="ortho",
                 activation="identity", use_grid=True):
        super().__init__()
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.hidden_dim = width
        self.n_layers = n_layers
        self.n_modes = n_modes
        self.grid = grid
        self.use_grid = grid  # reference quirk kept (ffno.py:51): the `use_grid` argument is overwritten by `grid`
        in_linear = self.in_channels + 1 if self.use_grid else self.in_channels
        self.in_proj = WNLinear(in_linear, self.hidden_dim, wnorm=ff_weight_norm)
        self.fourier_layers = nn.ModuleList([])
        for _ in range(n_layers):
            self.fourier_layers.append(FSpectralConv1d(self.hidden_dim, n_modes, factor=factor, ff_weight_norm=ff_weight_norm,
                                                       n_ff_layers=n_ff_layers, layer_norm=layer_norm, dropout=dropout,
                                                       mode=mode, fft_norm=fft_norm, activation=activation))
        self.out_proj = WNLinear(self.hidden_dim, self.out_channels, wnorm=ff_weight_norm)

    def forward(self, x):
        # ffno.py:98-125: (B, in, N) -> channels-last -> in_proj -> x += layer(x) -> out_proj -> (B, out, N)
        b, _, n = x.shape
        g = ops.coord_grid(self.grid, n, x.device) if self.use_grid else None
        x = ops.concat_grid(x, g, None, channels_last=True)          # (B, N, 1, in[+1])
        x = x.reshape(b, n, x.shape[-1])
        x = self.in_proj(x)
        for layer in self.fourier_layers:
            x, _ = layer(x, residual=x)                               # x + layer(x), skip fused into the layer
        return self.out_proj(x, channels_first_out=(b, self.out_channels, n))


class FFNO2D(nn.Module):
    def __init__(self, in_channels, out_channels, width=64, n_layers=4, n_modes=16, factor=4, ff_weight_norm=False,
                 n_ff_layers=2, layer_norm=False, grid=None, dropout=0.0, mode='full', use_grid=True):
        super().__init__()
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.hidden_dim = width
        self.n_layers = n_layers
        self.n_modes = n_modes
        self.ff_weight_norm = ff_weight_norm
        self.grid = grid
        self.use_grid = use_grid
        in_linear = self.in_channels + 2 if self.use_grid else self.in_channels
        if self.ff_weight_norm:
            self.in_proj = WNLinear(in_linear, self.hidden_dim, wnorm=ff_weight_norm)
        else:
            self.in_proj = Linear(in_linear, self.hidden_dim)
        self.fourier_layers = nn.ModuleList([])
        for _ in range(n_layers):
            self.fourier_layers.append(FSpectralConv2d(self.hidden_dim, n_modes, factor=factor, ff_weight_norm=ff_weight_norm,
                                                       n_ff_layers=n_ff_layers, layer_norm=layer_norm, dropout=dropout,
                                                       mode=mode))
        if self.ff_weight_norm:
            self.out_proj = WNLinear(self.hidden_dim, self.out_channels, wnorm=ff_weight_norm)
        else:
            self.out_proj = Linear(self.hidden_dim, self.out_channels)

    def forward(self, x):
        # ffno.py:210-237
        b, _, m, n = x.shape
        gx = gy = None
        if self.use_grid:
            gx = ops.coord_grid(self.grid[0] if self.grid is not None else None, m, x.device)
            gy = ops.coord_grid(self.grid[1] if self.grid is not None else None, n, x.device)
        x = ops.concat_grid(x, gx, gy, channels_last=True)           # (B, M, N, in[+2])
        x = self.in_proj(x)
        for layer in self.fourier_layers:
            x, _ = layer(x, residual=x)
        return self.out_proj(x, channels_first_out=(b, self.out_channels, m, n))
