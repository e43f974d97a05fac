"""ctypes binding of the C ABI in include/rpde_b200.h (librpde_b200.so).

The shared library is the product: it is built in-tree by ``build.py`` (nvcc, sm_100a) and loaded here.
There is no CPU or PyTorch fallback -- if the library is missing, or a tensor is not a CUDA tensor, the
call fails loudly.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_float, c_int, c_longlong, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librpde_b200.so")

P = c_void_p
I = c_int
L = c_longlong
Z = c_size_t
S = c_void_p  # stream

# name -> (restype, argtypes); mirrors include/rpde_b200.h one to one
SIGNATURES = {
    "rpde_version": (I, []),
    "rpde_last_error": (c_char_p, []),
    "rpde_set_precision": (I, [I]),
    "rpde_get_precision": (I, []),
    "rpde_launch_count": (ctypes.c_ulonglong, []),
    "rpde_tc_launch_count": (ctypes.c_ulonglong, []),
    "rpde_simt_launch_count": (ctypes.c_ulonglong, []),
    "rpde_trace_enable": (I, [I]),
    "rpde_trace_count": (I, []),
    "rpde_trace_get": (I, [I, P]),
    "rpde_dft_table": (I, [P, I, I, I, I, c_double, S]),
    "rpde_spectral1d_workspace_bytes": (Z, [I, I, I, I, I]),
    "rpde_spectral1d_fwd": (I, [P, P, P, P, P, P, P, I, I, I, I, I, I, I, P, Z, S]),
    "rpde_spectral1d_bwd": (I, [P, P, P, P, P, P, I, P, I, I, I, I, I, P, Z, S]),
    "rpde_fno_block1d_fwd": (I, [P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, P, Z, S]),
    "rpde_fno_block1d_bwd_fuses_dx": (I, []),
    "rpde_fno_block1d_bwd": (I, [P, P, I, P, P, P, P, P, P, P, I, P, I, I, I, I, I, P, Z, S]),
    "rpde_spectral2d_workspace_bytes": (Z, [I, I, I, I, I, I, I]),
    "rpde_spectral2d_fwd": (I, [P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, P, Z, S]),
    "rpde_spectral2d_bwd": (I, [P, P, P, P, P, P, P, P, P, I, P, P, I, I, I, I, I, I, I, P, Z, S]),
    "rpde_fno_block2d_supported": (I, [I, I, I, I, I, I, I]),
    "rpde_fno_block2d_workspace_bytes": (Z, [I, I, I, I, I, I, I]),
    "rpde_fno_block2d_fwd": (I, [P] * 12 + [I] * 8 + [P, Z, S]),
    "rpde_fno_block2d_bwd": (I, [P, P, I] + [P] * 12 + [I] * 7 + [P, Z, S]),
    "rpde_ffno_axis_workspace_bytes": (Z, [L, I, L, I, I]),
    "rpde_ffno_axis_fwd": (I, [P, P, P, P, P, P, I, L, I, L, I, I, I, I, P, Z, S]),
    "rpde_ffno_axis_bwd": (I, [P, P, P, P, P, P, I, P, I, L, I, L, I, I, I, I, P, Z, S]),
    "rpde_linear_fwd": (I, [P, P, P, P, P, L, I, I, I, L, S]),
    "rpde_linear_bwd_input": (I, [P, P, P, I, P, I, P, L, I, I, L, S]),
    "rpde_linear_bwd_params": (I, [P, P, P, P, L, I, I, L, S]),
    "rpde_conv1x1_fwd": (I, [P, P, P, P, P, I, I, I, L, I, I, S]),
    "rpde_conv1x1_bwd_input": (I, [P, P, P, I, P, I, I, I, I, L, S]),
    "rpde_conv1x1_bwd_params": (I, [P, P, P, P, I, I, I, L, S]),
    "rpde_mlp_supported": (I, [I, I, I, L]),
    "rpde_mlp_fwd": (I, [P, P, P, P, P, P, I, I, I, I, L, S]),
    "rpde_mlp_bwd": (I, [P, P, P, P, P, P, P, P, P, P, I, I, I, I, L, S]),
    "rpde_act_bwd": (I, [P, P, P, L, I, S]),
    "rpde_act_fwd": (I, [P, P, P, L, I, S]),
    "rpde_layernorm_fwd": (I, [P, P, P, P, P, P, P, L, I, c_float, S]),
    "rpde_layernorm_bwd": (I, [P, P, P, P, P, P, P, P, P, L, I, S]),
    "rpde_weight_norm_fwd": (I, [P, P, P, P, I, I, S]),
    "rpde_weight_norm_bwd": (I, [P, P, P, P, P, P, I, I, S]),
    "rpde_concat_grid_fwd": (I, [P, P, P, P, I, I, I, I, I, S]),
    "rpde_concat_grid_bwd": (I, [P, P, I, I, I, I, I, I, S]),
    "rpde_rel_l2_fwd": (I, [P, P, P, P, P, I, L, I, S]),
    "rpde_rel_l2_bwd": (I, [P, P, P, P, P, I, L, I, S]),
    "rpde_rel_l2_fwd_static": (I, [P, P, P, P, I, L, I, S]),
    "rpde_rollout_renorm": (I, [P, P, P, I, L, L, L, c_float, c_float, c_float, c_float, I, S]),
    "rpde_rollout_loss": (I, [P, P, P, P, P, I, I, L, L, L, L, L, c_float, c_float, I, S]),
    "rpde_moments_accumulate": (I, [P, L, P, S]),
    "rpde_moments_finalize": (I, [P, P, S]),
    "rpde_stride_downsample": (I, [P, P, L, I, I, I, I, I, S]),
    "rpde_binned_sums_f64": (I, [P, P, P, I, I, L, S]),
    "rpde_adamw_step": (I, [P, I, L, P, c_double, c_double, c_float, c_float, S]),
    "rpde_ff3_weight_image_bytes": (Z, []),
    "rpde_ff3_supported": (I, [I, I]),
    "rpde_ff3_pack_weights": (I, [P, P, P, P, S]),
    "rpde_ff3_fwd": (I, [P, P, P, P, P, P, P, P, P, L, c_float, S]),
    "rpde_ff3_bwd_workspace_bytes": (Z, [L]),
    "rpde_ff3_bwd": (I, [P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, L, c_float, P, Z, S]),
    "rpde_matmul": (I, [P, P, P, I, I, I, I, I, S]),
}

RPDE_E_MODES = -4
_lib = None


class ExtensionMissing(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    """Load librpde_b200.so (once).  Raises ExtensionMissing when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ExtensionMissing(
            f"{LIB_PATH} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()' "
            "or python resolution_pde_b200/build.py).  There is no CPU/PyTorch fallback for this path.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().rpde_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = "") -> None:
    """Translate a C-ABI return code into the exception the reference would raise."""
    if rc == 0:
        return
    msg = last_error()
    if rc == RPDE_E_MODES:
        # the reference fails inside torch.einsum with a RuntimeError (spectral_convolution.py:36,77)
        raise RuntimeError(msg)
    if rc == -1:
        raise ValueError(f"{what}: {msg}")
    raise RuntimeError(f"{what}: {msg} (code {rc})")
