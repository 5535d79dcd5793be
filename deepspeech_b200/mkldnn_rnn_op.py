"""API shape of the reference's src/mkldnn_rnn_op.py (a cell wrapper over the Intel-fork fused
`MkldnnRNNRelu` op with a flat parameter blob).  The native op is not available anywhere,
so parity is UNPINNED; the blob layout is build-defined: [W (H*I) | U (H*H) | b_W (H) | b_U (H)]
(cuDNN RNN_RELU order), plain ReLU, no batch norm, no clipping."""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from .custom_ops import _AddFn, _get_variable


class _MkldnnRNNRelu(object):
    """Stand-in for mkldnn_rnn_ops.MkldnnRNNRelu(num_layers, num_units, input_size, dropout)."""

    def __init__(self, num_layers, num_units, input_size, dropout=0.0):
        if num_layers != 1 or dropout != 0.0:
            raise ValueError("MkldnnRNNRelu: the reference only builds num_layers=1, dropout=0.0")
        self.num_units, self.input_size = num_units, input_size

    def params_size(self):
        H, I = self.num_units, self.input_size
        return H * I + H * H + 2 * H

    def __call__(self, input_data, input_h, params, is_training=True):
        H, I = self.num_units, self.input_size
        W = params[:H * I].view(H, I)
        U = params[H * I:H * I + H * H].view(H, H)
        bW = params[H * I + H * H:H * I + H * H + H]
        bU = params[H * I + H * H + H:]
        out = ops.relux(_AddFn.apply(ops.linear(input_data.float(), W, bW, False), ops.linear(input_h.float(), U, bU, False)), None)
        return out, out

    def sequence(self, input_data, seq_lengths, params, reverse=False, use_bf16=False):
        """The whole-sequence form the Intel op really is (src/mkldnn_rnn_op.py:22-29 builds a num_layers=1 sequence op and
        the reference then calls it once per timestep): input_data [T,B,I] time-major, seq_lengths int32 [B] (device),
        params the flat blob -> outputs [T,B,H].  One batched W.x GEMM + the recurrence kernels (persistent tcgen05 kernel on
        the bf16 path), plain ReLU (no clip, no seq-BN scale), bias = b_W + b_U; gradients reach the blob through its views."""
        H, I = self.num_units, self.input_size
        W = params[:H * I].view(H, I)
        U = params[H * I:H * I + H * H].view(1, H, H)
        bias = (params[H * I + H * H:H * I + H * H + H] + params[H * I + H * H + H:]).view(1, H)
        odt = torch.bfloat16 if use_bf16 else torch.float32
        return ops.birnn_layer(input_data, W, U, bias, seq_lengths, 1 if reverse else 0, use_bf16, odt, scale=1.0, cap=3.0e38)


class MkldnnRNNCell(object):
    """src/mkldnn_rnn_op.py:16-45.  `sess` is accepted and ignored."""

    def __init__(self, sess, num_units, input_size=None, activation=None, use_fp16=False):
        self._num_units = num_units
        self.use_fp16 = use_fp16
        self.model = _MkldnnRNNRelu(1, num_units, input_size, dropout=0.0)
        self.param_size = self.model.params_size()

    @property
    def state_size(self):
        return self._num_units

    @property
    def output_size(self):
        return self._num_units

    def zero_state(self, batch_size, dtype=torch.float32, device="cuda"):
        return torch.zeros(batch_size, self._num_units, dtype=dtype, device=device)

    def weights(self, scope=None, device="cuda"):
        sc = scope or type(self).__name__
        return _get_variable(sc, "rnn_weights", (self.param_size,), 1.0 / self.param_size, True, device)          # :39

    def __call__(self, inputs, state, scope=None, weight_size=None):
        w = self.weights(scope, inputs.device)
        return self.model(input_data=inputs, input_h=state, params=w, is_training=True)

    def sequence(self, inputs, seq_lengths, scope=None, reverse=False):
        """All timesteps at once through the sequence kernels, from the same flat blob the per-step call uses."""
        return self.model.sequence(inputs, seq_lengths, self.weights(scope, inputs.device), reverse, self.use_fp16)
