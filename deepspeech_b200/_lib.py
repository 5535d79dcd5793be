"""ctypes binding of libds2_b200.so (the C ABI declared in include/ds2_b200.h).

The product has exactly one compute backend: the hand-written sm_100a kernels in csrc/.  If the
shared library is missing this module raises at import - there is no CPU or library fallback.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_float, c_int, c_long, c_longlong, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libds2_b200.so")

DS2_F32, DS2_BF16 = 0, 1
_ERR_NAMES = {1: "BAD_ARG", 2: "BAD_SHAPE", 3: "WORKSPACE", 4: "UNSUPPORTED", 5: "CUDA"}

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "deepspeech_b200: %s not found - build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "or deepspeech_b200/csrc/build.sh (there is no CPU fallback)" % LIB_PATH)

lib = ctypes.CDLL(LIB_PATH)

P, I, L, F, D, SZ = c_void_p, c_int, c_long, c_float, c_double, c_size_t
_SIGS = {
    "ds2_last_error": (c_char_p, []),
    "ds2_version": (I, []),
    "ds2_launch_count": (c_longlong, []),
    "ds2_debug_reload_env": (None, []),
    "ds2_ctc_workspace_bytes": (SZ, [I, I, I, I]),
    "ds2_ctc_loss_fwd_bwd": (I, [P, I, I, I, P, I, P, P, P, P, P, P, SZ, P]),
    "ds2_ctc_noalloc_max_bytes": (SZ, []),
    "ds2_ctc_loss_fwd_bwd_noalloc": (I, [P, I, I, I, P, I, P, P, P, P, P, P]),
    "ds2_ctc_scale_grad": (I, [P, P, P, I, I, I, P]),
    "ds2_gemm_bf16": (I, [P, L, I, P, L, I, P, I, L, I, I, I, F, F, P, P]),
    "ds2_gemm_bf16_splitk": (I, [P, L, I, P, L, I, P, L, I, I, I, F, P, I, P]),
    "ds2_im2col": (I, [P, P, P, P, P, F, I, P, L, I, I, I, I, I, I, I, I, P]),
    "ds2_col2im": (I, [P, L, P, I, I, I, I, I, I, I, I, P]),
    "ds2_gemm_generic": (I, [P, I, L, L, P, I, L, L, P, I, L, I, I, I, F, F, P, P]),
    "ds2_rnn_relu_fwd": (I, [P, F, P, P, P, P, I, I, I, I, I, I, F, P, SZ, P]),
    "ds2_rnn_relu_fwd_ex": (I, [P, F, P, P, P, P, P, I, I, I, I, I, F, P, SZ, P]),
    "ds2_rnn_relu_bwd_ex": (I, [P, P, P, P, P, P, P, I, I, I, I, I, F, P, SZ, P]),
    "ds2_rnn_relu_bwd": (I, [P, P, P, P, P, P, I, I, I, I, I, I, F, P, SZ, P]),
    "ds2_rnn_workspace_bytes": (SZ, [I, I, I, I, I]),
    "ds2_rnn_relu_persistent_supported": (I, [I, I, I]),
    "ds2_gru_fwd": (I, [P, F, P, P, P, P, P, P, I, I, I, I, I, I, P]),
    "ds2_gru_bwd": (I, [P, P, P, P, P, P, P, P, I, I, I, I, I, I, P]),
    "ds2_set_sm_limit": (I, [I]),
    "ds2_ctc_greedy_decode": (I, [P, I, I, I, P, P, P, P]),
    "ds2_ctc_labels_from_sparse": (I, [P, P, I, I, I, I, I, P, P, P, P, P]),
    "ds2_dummy_sparse_labels": (I, [P, P, I, I, P, P, P]),
    "ds2_edit_distance": (I, [P, I, P, P, I, P, I, I, P, P]),
    "ds2_adam_ema_chunk": (I, []),
    "ds2_adam_ema_step": (I, [P, I, I, F, F, F, F, L, F, F, P]),
    "ds2_gru_persistent_supported": (I, [I, I, I]),
    "ds2_gru_workspace_bytes": (SZ, [I, I, I, I]),
    "ds2_gru_persistent_fwd": (I, [P, F, P, P, P, P, P, P, I, I, I, I, I, P, SZ, P]),
    "ds2_gru_persistent_bwd": (I, [P, P, P, P, P, P, P, P, P, I, I, I, I, I, P, SZ, P]),
    "ds2_conv2d_fwd": (I, [P, I, P, P, P, I, I, I, I, I, I, I, I, I, P]),
    "ds2_conv2d_dgrad": (I, [P, P, P, I, I, I, I, I, I, I, I, I, P]),
    "ds2_conv2d_wgrad": (I, [P, I, P, P, P, I, I, I, I, I, I, I, I, I, P]),
    "ds2_conv1_fused_supported": (I, [I, I, I, I, I, I, I, I]),
    "ds2_conv1_fused_fwd": (I, [P, P, P, P, P, I, I, I, I, I, I, I, P]),
    "ds2_conv1_fused_wgrad": (I, [P, P, P, P, I, I, I, I, I, I, I, P]),
    "ds2_bn_finalize": (I, [P, D, I, F, P, P, P]),
    "ds2_conv_window_supported": (I, [I, I, I, I, I, I]),
    "ds2_conv_window_weights": (I, [P, P, P, I, I, P]),
    "ds2_conv_window_dy_layout": (I, [I, I, I, I, I, P, P]),
    "ds2_conv_window_pad_cast": (I, [P, P, I, I, I, I, I, I, P]),
    "ds2_conv_window_fwd": (I, [P, P, P, P, P, I, I, I, I, I, I, P]),
    "ds2_conv_window_dgrad": (I, [P, P, P, I, I, I, I, I, I, P]),
    "ds2_conv_window_wgrad": (I, [P, P, P, I, I, I, I, I, I, P]),
    "ds2_bn_stats": (I, [P, L, I, F, P, P, P, P]),
    "ds2_bn_clip_fwd": (I, [P, I, I, I, I, P, P, P, P, F, I, P, I, L, L, L, L, P]),
    "ds2_bn_clip_bwd": (I, [P, I, I, I, I, P, P, P, P, F, I, P, I, L, L, L, L, P, P, P, P, P]),
    "ds2_relux_fwd": (I, [P, P, L, F, I, P]),
    "ds2_relux_bwd": (I, [P, P, P, L, F, I, P]),
    "ds2_scale": (I, [P, F, P, L, P]),
    "ds2_add2": (I, [P, P, I, P, I, L, P]),
    "ds2_cast": (I, [P, I, P, I, L, P]),
    "ds2_colsum": (I, [P, I, L, I, L, P, I, P]),
    "ds2_transpose": (I, [P, I, L, L, L, P, I, L, P]),
}
EXPORTS = sorted(_SIGS)
for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)          # AttributeError here = header/library mismatch: fail loudly
    _fn.restype = _res
    _fn.argtypes = _args


def last_error() -> str:
    return (lib.ds2_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    """Map C status codes to the Python exceptions the reference's call sites would see:
    bad shapes/arguments -> ValueError (graph-build errors, src/custom_ops.py:219-233), the rest ->
    RuntimeError (TF run-time InvalidArgumentError analogue)."""
    if rc == 0:
        return
    msg = "%s failed [%s]: %s" % (what, _ERR_NAMES.get(rc, rc), last_error())
    if rc in (1, 2):
        raise ValueError(msg)
    raise RuntimeError(msg)
