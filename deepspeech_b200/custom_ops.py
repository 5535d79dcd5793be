"""Operator surface of the reference's src/custom_ops.py, same names and argument lists, on CUDA torch
tensors.  Weights that TF kept in variable scopes live in each op object's `variables` dict, keyed by
the scope string, so `[cell] * L` shares the object but not the weights exactly like the reference
(SURVEY Q4)."""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from .helper_routines import glorot_uniform, stable_seed

_SCOPE_VARS = {}      # (scope-name) -> dict of Parameters; the analogue of TF's variable store


def variable_store():
    return _SCOPE_VARS


def reset_variable_store():
    _SCOPE_VARS.clear()


def _get_variable(scope, name, shape, init, trainable=True, device="cuda"):
    key = "%s/%s" % (scope, name)
    v = _SCOPE_VARS.get(key)
    if v is None:
        val = init(shape) if callable(init) else np.full(shape, float(init), dtype=np.float32)
        v = torch.nn.Parameter(torch.from_numpy(np.asarray(val, dtype=np.float32)).to(device), requires_grad=trainable)
        _SCOPE_VARS[key] = v
    elif tuple(v.shape) != tuple(shape):
        raise ValueError("variable %s exists with shape %s, requested %s" % (key, tuple(v.shape), tuple(shape)))
    return v


def relux(x, capping=None):
    """Clipped ReLU, src/custom_ops.py:99-104.  (capping=None raises NameError in the reference; here it
    is plain ReLU.)"""
    return ops.relux(x, capping)


def _to_channel_last(x, data_format):
    return x.permute(0, 2, 3, 1).contiguous() if data_format == "NCHW" else x.contiguous()


def _from_channel_last(y, data_format):
    return y.permute(0, 3, 1, 2).contiguous() if data_format == "NCHW" else y


def batch_norm2(inputs, decay=0.999, center=True, scale=True, epsilon=0.001, moving_vars="moving_vars", activation=None,
                is_training=True, trainable=True, scope=None, reuse=None, data_format="NHWC"):
    """src/custom_ops.py:107-160: tf.nn.fused_batch_norm in training mode (batch mean, biased batch
    variance, eps 1e-3), beta/gamma created in variable scope 'bn2'."""
    if not is_training:
        raise NotImplementedError("only training-mode BN is on the reference's hot path")
    x = _to_channel_last(inputs, data_format)
    C = x.shape[-1]
    sc = "%s/bn2" % (scope or "")
    beta = _get_variable(sc, "beta", (C,), 0.0, trainable and center, x.device)
    gamma = _get_variable(sc, "gamma", (C,), 1.0, trainable and scale, x.device)
    y = ops.bn_clip_nhwc(x, gamma, beta, float(epsilon), None)
    return _from_channel_last(y, data_format)


def batch_norm(x, scope=None, is_train=True, data_format=None):
    """src/custom_ops.py:163-181: moments over [0,1,2] + batch_normalization, eps 1e-5, NHWC only.
    (is_train=False is broken in the reference)"""
    if not is_train:
        raise NotImplementedError("is_train=False is broken in the reference (custom_ops.py:176-179)")
    C = x.shape[-1]
    sc = scope or "bn"
    beta = _get_variable(sc, "beta", (C,), 0.0, True, x.device)
    gamma = _get_variable(sc, "gamma", (C,), 1.0, True, x.device)
    return ops.bn_clip_nhwc(x.contiguous(), gamma, beta, 1e-5, None)


def seq_batch_norm(x, scope=None, is_train=True):
    """src/custom_ops.py:184-199 as written: the moving statistics are never updated, so this is
    (x - 0) / sqrt(1 + 1e-5) * 1 + 0.  Variables are created (non-trainable) for parity of
    the variable list."""
    D = x.shape[-1]
    for name, val in (("beta", 0.0), ("gamma", 1.0), ("moving_mean", 0.0), ("moving_variance", 1.0)):
        _get_variable("sbn/%d" % D, name, (D,), val, False, x.device)
    return _ScaleFn.apply(x, ops.SBN_SCALE)


class _ScaleFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, s):
        ctx.s = s
        xc = x.contiguous()
        out = torch.empty_like(xc)
        ops.check(ops.lib.ds2_scale(ops._p(xc), float(s), ops._p(out), xc.numel(), ops._stream()), "ds2_scale")
        return out

    @staticmethod
    def backward(ctx, g):
        return _ScaleFn.apply(g, ctx.s), None


def _linear(args, output_size, bias, scope=None, use_fp16=False):
    """src/custom_ops.py:202-257: sum_i args[i] . Matrix^T (+ Bias); same ValueErrors."""
    if args is None or (isinstance(args, (list, tuple)) and not args):
        raise ValueError("`args` must be specified")
    if not isinstance(args, (list, tuple)):
        args = [args]
    total = 0
    shapes = [list(a.shape) for a in args]
    for shape in shapes:
        if len(shape) != 2:
            raise ValueError("Linear is expecting 2D arguments: %s" % str(shapes))
        if not shape[1]:
            raise ValueError("Linear expects shape[1] of arguments: %s" % str(shapes))
        total += shape[1]
    sc = scope or "Linear"
    # the reference declares Matrix [total, output] and multiplies with transpose_b=True (SURVEY a13:
    # shape-consistent only as [output, total]); the build uses the consistent shape.
    matrix = _get_variable(sc, "Matrix", (output_size, total), lambda s: glorot_uniform(np.random.default_rng(0), s),
                           True, args[0].device)
    x = args[0] if len(args) == 1 else torch.cat(list(args), 1)
    b = _get_variable(sc, "Bias", (output_size,), 0.0, True, x.device) if bias else None
    return ops.linear(x.float(), matrix, b, False)


class CustomRNNCell2(object):
    """src/custom_ops.py:36-72: h' = min(relu(seq_batch_norm(x W^T) + h U^T + B), 20); returns (h', h')."""

    def __init__(self, num_units, input_size=None, activation=None, use_fp16=False):
        self._num_units = num_units
        self.use_fp16 = use_fp16

    @property
    def state_size(self):
        return self._num_units

    @property
    def output_size(self):
        return self._num_units

    def zero_state(self, batch_size, dtype=torch.float32, device="cuda"):
        return torch.zeros(batch_size, self._num_units, dtype=dtype, device=device)

    def variables(self, scope, input_size, device="cuda"):
        sc = "%s/CustomRNNCell2" % scope if scope else "CustomRNNCell2"
        H = self._num_units
        seed = stable_seed(sc)
        W = _get_variable(sc, "W", (H, input_size), lambda s: glorot_uniform(np.random.default_rng(seed), s), True, device)
        U = _get_variable(sc, "U", (H, H), lambda s: glorot_uniform(np.random.default_rng(seed + 1), s), True, device)
        B = _get_variable(sc, "B", (H,), 0.0, True, device)
        return W, U, B

    def __call__(self, inputs, state, scope=None):
        W, U, B = self.variables(scope, inputs.shape[1], inputs.device)
        resi = seq_batch_norm(ops.linear(inputs.float(), W, None, False))
        resu = ops.linear(state.float(), U, B, False)
        out = relux(_AddFn.apply(resi, resu), capping=20)
        return out, out


class CustomRNNCell(CustomRNNCell2):
    """src/custom_ops.py:14-33: relu6(_linear([x, h])).  Dead code in the reference (never instantiated,
    SURVEY a13); kept for surface completeness."""

    def __call__(self, inputs, state, scope=None):
        out = relux(_linear([inputs, state], self._num_units, True, scope=scope or type(self).__name__), capping=6)
        return out, out


class _AddFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        return ops.add2(a.contiguous(), b.contiguous(), a.dtype)

    @staticmethod
    def backward(ctx, g):
        return g, g


def stacked_brnn(cell_fw, cell_bw, num_units, num_layers, inputs, seq_lengths, batch_size, wiring="as_written",
                 scope="rnn"):
    """src/custom_ops.py:75-96.  inputs [T,B,I] time-major, seq_lengths int [B] -> [T,B,num_units].

    wiring='as_written' reproduces the reference verbatim: every layer reads the ORIGINAL inputs and only
    the last layer's fw+bw is returned (`inputs = inputs` / `_inputs = ...`, lines 86,95) - the
    dead layers are not executed, as TF prunes them.  wiring='stacked' is the docstring's intent."""
    if wiring not in ("as_written", "stacked"):
        raise ValueError("stacked_brnn: wiring must be 'as_written' or 'stacked'")
    T, B, I = inputs.shape
    if B != batch_size:
        raise ValueError("stacked_brnn: batch_size %d != inputs.shape[1] %d" % (batch_size, B))
    lens_np = np.minimum(np.asarray(seq_lengths.detach().cpu().numpy() if isinstance(seq_lengths, torch.Tensor)
                                    else seq_lengths, dtype=np.int32), T)
    lens = torch.from_numpy(lens_np).to(inputs.device)
    want_bf16 = bool(getattr(cell_fw[0], "use_fp16", False))
    act = torch.bfloat16 if want_bf16 else torch.float32
    cur = inputs
    out = None
    for i in range(num_layers):
        live = (wiring == "stacked") or (i == num_layers - 1)
        Wf, Uf, Bf = cell_fw[i].variables("%s/brnn-%d/bidirectional_rnn/fw" % (scope, i), cur.shape[2], inputs.device)
        Wb, Ub, Bb = cell_bw[i].variables("%s/brnn-%d/bidirectional_rnn/bw" % (scope, i), cur.shape[2], inputs.device)
        if not live:
            continue
        out = ops.birnn_layer(cur, torch.cat([Wf, Wb], 0), torch.stack([Uf, Ub], 0), torch.stack([Bf, Bb], 0), lens, 0b10,
                              want_bf16, act)
        if wiring == "stacked":
            cur = out
    return out
