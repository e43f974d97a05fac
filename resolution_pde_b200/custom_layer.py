"""FeedForward and WNLinear of the F-FNO (reference: models/custom_layer.py:49-108), same constructor
arguments and state-dict layout, arithmetic in librpde_b200.so."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops


class Linear(nn.Linear):
    """nn.Linear whose forward/backward run in the CUDA extension (channels-last, any leading dims)."""

    def forward(self, x, channels_first_out=None):
        return ops.linear(x, self.weight, self.bias, channels_first_out)


class WNLinear(nn.Linear):
    """custom_layer.py:70-108.  With wnorm=True the parameters are `weight_g` (out,1) and `weight_v`
    (out,in) exactly as torch.nn.utils.weight_norm (dim=0) registers them, so reference checkpoints load
    unchanged; w = g * v / ||v||_row is rebuilt by a CUDA kernel on every forward."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, device=None, dtype=None, wnorm=False):
        super().__init__(in_features=in_features, out_features=out_features, bias=bias, device=device, dtype=dtype)
        self.wnorm = bool(wnorm)
        if self.wnorm:
            w = self.weight
            del self._parameters["weight"]
            # same initial values as torch's weight_norm: g = row norms, v = the original weight
            self.register_parameter("weight_g", nn.Parameter(w.detach().norm(2, dim=1, keepdim=True)))
            self.register_parameter("weight_v", nn.Parameter(w.detach().clone()))

    def effective_weight(self):
        if self.wnorm:
            return ops.weight_norm(self.weight_v, self.weight_g)
        return self.weight

    def forward(self, x, channels_first_out=None):
        return ops.linear(x, self.effective_weight(), self.bias, channels_first_out)


class FeedForward(nn.Module):
    """custom_layer.py:49-68.  `ff_weight_norm` is accepted and ignored, as in the reference.  Layer i is
    Sequential(Linear, Dropout, GELU|Identity, LayerNorm|Identity) so the state-dict keys are
    layers.{i}.0.{weight,bias} and layers.{n-1}.3.{weight,bias}."""

    def __init__(self, dim, factor, n_layers=2, ff_weight_norm=False, layer_norm=False, dropout=0.0):
        super().__init__()
        self.layers = nn.ModuleList([])
        self.n_layers = n_layers
        self.use_layer_norm = bool(layer_norm)
        self.dropout = dropout
        for i in range(n_layers):
            in_dim = dim if i == 0 else dim * factor
            out_dim = dim if i == n_layers - 1 else dim * factor
            self.layers.append(nn.Sequential(
                Linear(in_dim, out_dim),
                nn.Dropout(dropout),
                nn.GELU() if i < n_layers - 1 else nn.Identity(),
                nn.LayerNorm(out_dim) if layer_norm and i == n_layers - 1 else nn.Identity(),
            ))

    def forward(self, x, residual=None, final_act: int = ops.ACT_NONE):
        """y = residual + act(FeedForward(x)); the two extras let the F-FNO layer fuse its activation and the
        model's skip connection into the same kernels (spectral_convolution.py:155; ffno.py:116-118, 228-230)."""
        if self.training and self.dropout > 0.0:
            return self._forward_dropout(x, residual, final_act)
        ws = [l[0].weight for l in self.layers]
        bs = [l[0].bias for l in self.layers]
        ln = self.layers[-1][3] if self.use_layer_norm else None
        return ops.feed_forward(x, residual, ws, bs, ln.weight if ln is not None else None,
                                ln.bias if ln is not None else None, final_act, ln.eps if ln is not None else 1e-5)

    def _forward_dropout(self, x, residual, final_act):
        # Train-mode dropout (p > 0) draws torch's Philox mask between our kernels.  Parity with the
        # reference is defined at p = 0 (the mask cannot match bit for bit); this path keeps semantics.
        n = self.n_layers
        for i, l in enumerate(self.layers):
            last = i == n - 1
            if not last:
                x = ops.feed_forward(l[1](ops.linear(x, l[0].weight, l[0].bias)), None, [], [], None, None, ops.ACT_GELU)
            else:
                x = l[1](ops.linear(x, l[0].weight, l[0].bias))
                ln = l[3] if self.use_layer_norm else None
                x = ops.feed_forward(x, residual, [], [], ln.weight if ln is not None else None,
                                     ln.bias if ln is not None else None, final_act, ln.eps if ln is not None else 1e-5)
        return x
