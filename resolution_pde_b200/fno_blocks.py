"""FNO blocks and projection MLPs (reference: models/fno_blocks.py:25-83)."""
from __future__ import annotations

import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .spectral_convolution import SpectralConv1d, SpectralConv2d


class Conv1d1x1(nn.Conv1d):
    """nn.Conv1d(kernel_size=1) with the reference's parameter shapes, run by the CUDA extension."""

    def forward(self, x):
        return ops.conv1x1(x, self.weight, self.bias)


class Conv2d1x1(nn.Conv2d):
    def forward(self, x):
        return ops.conv1x1(x, self.weight, self.bias)


def _block_forward(self, x):
    aid = ops.act_id(self.activation)
    sc = self.spectral_conv
    w2 = getattr(sc, "weights2", None)
    if aid is None:  # unknown callable: fuse everything but the activation
        return self.activation(ops.fno_spectral(x, sc.weights1, w2, self.bypass_conv.weight, self.bypass_conv.bias))
    return ops.fno_spectral(x, sc.weights1, w2, self.bypass_conv.weight, self.bypass_conv.bias, aid)


class FNOBlock1d(nn.Module):
    """fno_blocks.py:25-33: activation(spectral_conv(x) + bypass_conv(x)) as one fused operator."""

    def __init__(self, in_channels, out_channels, modes, activation=F.relu):
        super().__init__()
        self.spectral_conv = SpectralConv1d(in_channels, out_channels, modes)
        self.bypass_conv = Conv1d1x1(in_channels, out_channels, 1)
        self.activation = activation

    forward = _block_forward


class MLP1d(nn.Module):
    """fno_blocks.py:35-45: conv1x1 -> GELU -> conv1x1."""

    def __init__(self, in_channels, out_channels, mid_channels):
        super().__init__()
        self.mlp1 = Conv1d1x1(in_channels, mid_channels, 1)
        self.mlp2 = Conv1d1x1(mid_channels, out_channels, 1)

    def forward(self, x):
        return ops.mlp_gelu(x, self.mlp1.weight, self.mlp1.bias, self.mlp2.weight, self.mlp2.bias)


class FNOBlock2d(nn.Module):
    """fno_blocks.py:63-71."""

    def __init__(self, in_channels, out_channels, modes1, modes2, activation=F.gelu):
        super().__init__()
        self.spectral_conv = SpectralConv2d(in_channels, out_channels, modes1, modes2)
        self.bypass_conv = Conv2d1x1(in_channels, out_channels, 1)
        self.activation = activation

    forward = _block_forward


class MLP2d(nn.Module):
    """fno_blocks.py:73-83."""

    def __init__(self, in_channels, out_channels, mid_channels):
        super().__init__()
        self.mlp1 = Conv2d1x1(in_channels, mid_channels, 1)
        self.mlp2 = Conv2d1x1(mid_channels, out_channels, 1)

    def forward(self, x):
        return ops.mlp_gelu(x, self.mlp1.weight, self.mlp1.bias, self.mlp2.weight, self.mlp2.bias)
