"""resolution-pde, B200-native spectral path.

Drop-in replacements for the reference's spectral layers and the two model families built on them
(models/spectral_convolution.py, models/fno.py, models/fno_blocks.py, models/ffno.py,
models/custom_layer.py: FeedForward/WNLinear, utils/loss.py), backed by hand-written sm_100a CUDA kernels
behind the C ABI in include/rpde_b200.h.
"""
from . import _lib, ops
from .custom_layer import FeedForward, Linear, WNLinear
from .ffno import FFNO1D, FFNO2D
from .fno import FNO1d, FNO2d
from .fno_blocks import FNOBlock1d, FNOBlock2d, MLP1d, MLP2d
from .loss import RelativeL2Loss
from .ops import get_precision, set_precision
from .training import GraphedTrainStep, ResolutionGroupedBatches
from .optim import FusedAdamW
from .rollout import RolloutEvaluator, SimpleNormalizer, evaluate_rollout, perform_rollout_1d, perform_rollout_2d, rollout_losses
from .spectral_convolution import FSpectralConv1d, FSpectralConv2d, SpectralConv1d, SpectralConv2d

__all__ = [
    "SpectralConv1d", "SpectralConv2d", "FSpectralConv1d", "FSpectralConv2d", "FNOBlock1d", "FNOBlock2d", "MLP1d",
    "MLP2d", "FNO1d", "FNO2d", "FFNO1D", "FFNO2D", "FeedForward", "WNLinear", "Linear", "RelativeL2Loss",
    "set_precision", "get_precision", "ops", "GraphedTrainStep", "ResolutionGroupedBatches", "FusedAdamW", "RolloutEvaluator", "SimpleNormalizer",
    "evaluate_rollout", "perform_rollout_1d", "perform_rollout_2d", "rollout_losses",
]
