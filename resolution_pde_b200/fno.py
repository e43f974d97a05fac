"""FNO1d / FNO2d (reference: models/fno.py:24-150): same constructors, forward signatures and state-dict."""
from __future__ import annotations

import math

import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .fno_blocks import Conv1d1x1, Conv2d1x1, FNOBlock1d, FNOBlock2d, MLP1d, MLP2d


class FNO1d(nn.Module):
    def __init__(self, in_channels, out_channels, modes, width, grid=None, activation=F.relu, n_blocks=4):
        super().__init__()
        self.width = width
        self.grid = grid
        self.lifting = Conv1d1x1(in_channels + 1, width, 1)  # 1 for the grid coordinate
        self.fno_blocks = nn.ModuleList([FNOBlock1d(width, width, modes, activation) for _ in range(n_blocks)])
        self.projection = MLP1d(self.width, out_channels, self.width * 4)

    def forward(self, x):
        # fno.py:56-76.  The coordinate channel (linspace(0, 2*pi, N) unless a grid was given) is cached on the
        # device per resolution instead of being rebuilt on the host and copied every call.
        n = x.shape[-1]
        g = ops.coord_grid(self.grid, n, x.device, 0.0, 2 * math.pi)
        x = ops.concat_grid(x, g, None, channels_last=False)
        x = self.lifting(x)
        for blk in self.fno_blocks:
            x = blk(x)
        return self.projection(x)


class FNO2d(nn.Module):
    def __init__(self, in_channels, out_channels, modes1, modes2, width, grid=None, activation=F.gelu, n_blocks=4):
        super().__init__()
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.modes1 = modes1
        self.modes2 = modes2
        self.width = width
        self.grid = grid
        self.lifting = Conv2d1x1(in_channels + 2, width, 1)  # 2 for the grid coordinates
        self.fno_blocks = nn.ModuleList([FNOBlock2d(width, width, modes1, modes2, activation) for _ in range(n_blocks)])
        self.projection = MLP2d(self.width, self.out_channels, self.width * 4)

    def forward(self, x):
        # fno.py:131-150; grid = linspace(0, 1) per axis
        h, w = x.shape[-2], x.shape[-1]
        gx = ops.coord_grid(self.grid[0] if self.grid is not None else None, h, x.device)
        gy = ops.coord_grid(self.grid[1] if self.grid is not None else None, w, x.device)
        x = ops.concat_grid(x, gx, gy, channels_last=False)
        x = self.lifting(x)
        for blk in self.fno_blocks:
            x = blk(x)
        return self.projection(x)
