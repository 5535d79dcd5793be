"""DeepSpeech2 model of the reference (conv x2 -> stacked bi-RNN -> FC), parameters as torch
Parameters, compute through the ds2_b200 kernels.  Mirrors `inference()` of
src/deepSpeech_NCHW.py:81-201 (layout='NCHW') and src/deepSpeech.py:81-203 (layout='NHWC')."""
from __future__ import annotations

import math
from typing import Dict, List

import numpy as np
import torch

from . import ops
from .helper_routines import glorot_uniform, variance_scaling

NUM_CLASSES = 29            # src/deepSpeech_input.py:13-14
FREQ_BINS = 161             # src/deepSpeech_dummy.py:13


def get_rnn_seqlen(seq_lens):
    """src/deepSpeech.py:38-48 (float64 div/ceil, int32 result); host-side integer math."""
    s = np.asarray(seq_lens.detach().cpu().numpy() if isinstance(seq_lens, torch.Tensor) else seq_lens, dtype=np.float64)
    s = np.ceil((s - 19.0) / 2.0)
    s = np.ceil((s - 9.0) / 2.0)
    return s.astype(np.int32)


class DeepSpeech2(torch.nn.Module):
    def __init__(self, num_filters=32, num_hidden=1760, num_rnn_layers=7, layout="NCHW", wiring="stacked",
                 cell="relu_clip", use_bf16=False, rnn_type="bi-dir", seed=1234, device="cuda", engine="mkl"):
        super().__init__()
        if wiring not in ("as_written", "stacked", "two_stack"):
            raise ValueError("wiring must be as_written | stacked | two_stack")
        if cell not in ("relu_clip", "gru"):
            raise ValueError("cell must be relu_clip | gru")
        self.nf, self.H, self.L = num_filters, num_hidden, num_rnn_layers
        self.layout, self.wiring, self.cell, self.use_bf16 = layout, wiring, cell, use_bf16
        # engine 'mkldnn_rnn' / 'cudnn_rnn' (src/deepSpeech_NCHW.py:173-176): the recurrent layers are MkldnnRNNRelu cells -
        # one FLAT parameter blob per (layer, direction), plain ReLU, no seq-BN, no clip (src/mkldnn_rnn_op.py:22-45)
        self.engine = engine
        self.blob_engine = engine in ("mkldnn_rnn", "cudnn_rnn") and cell == "relu_clip" and wiring in ("as_written", "stacked")
        self.bidirectional = rnn_type != "uni-dir"          # src/deepSpeech.py:173
        self.eps = 1e-3 if layout == "NCHW" else 1e-5       # batch_norm2 vs batch_norm (custom_ops.py:111,180)
        rng = np.random.default_rng(seed)
        nf, H = num_filters, num_hidden
        P = lambda a: torch.nn.Parameter(torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(device))
        self.conv1_w = P(variance_scaling(rng, (20, 5, 1, nf)))
        self.conv1_b = P(np.full(nf, -0.05))
        self.conv1_gamma, self.conv1_beta = P(np.ones(nf)), P(np.zeros(nf))
        self.conv2_w = P(variance_scaling(rng, (10, 5, nf, nf)))
        self.conv2_b = P(np.full(nf, -0.05))
        self.conv2_gamma, self.conv2_beta = P(np.ones(nf)), P(np.zeros(nf))
        self.fc_w = P(variance_scaling(rng, (NUM_CLASSES, H)))
        self.fc_b = P(np.zeros(NUM_CLASSES))
        G = 3 if cell == "gru" else 1
        self.rnn_W = torch.nn.ParameterList()
        self.rnn_U = torch.nn.ParameterList()
        self.rnn_B = torch.nn.ParameterList()      # relu: [2,H] ; gru: bw [2,3H]
        self.rnn_BU = torch.nn.ParameterList()     # gru only
        self.rnn_blob = torch.nn.ParameterList()   # blob engines: [fw_0, bw_0, fw_1, bw_1, ...]
        for l in range(num_rnn_layers):
            I = 75 * nf if (l == 0 or wiring == "as_written") else H
            if self.blob_engine:
                n = H * I + H * H + 2 * H
                for _ in range(2):
                    self.rnn_blob.append(P(np.full(n, 1.0 / n)))            # src/mkldnn_rnn_op.py:39: constant 1/param_size
                continue
            self.rnn_W.append(P(np.concatenate([glorot_uniform(rng, (G * H, I)) for _ in range(2)], 0)))
            self.rnn_U.append(P(np.stack([glorot_uniform(rng, (G * H, H)) for _ in range(2)], 0)))
            self.rnn_B.append(P(np.zeros((2, G * H))))
            if cell == "gru":
                self.rnn_BU.append(P(np.zeros((2, G * H))))

    # ---- weights from the oracle's dict format (tests / bench use identical weights on both sides)
    @torch.no_grad()
    def load_oracle_params(self, p: Dict[str, torch.Tensor], layers: List[dict]):
        def put(dst, src):
            dst.copy_(torch.as_tensor(np.asarray(src), dtype=torch.float32).reshape(dst.shape))
        put(self.conv1_w, p["conv1.weights"]); put(self.conv1_b, p["conv1.biases"])
        put(self.conv1_gamma, p["conv1.gamma"]); put(self.conv1_beta, p["conv1.beta"])
        put(self.conv2_w, p["conv2.weights"]); put(self.conv2_b, p["conv2.biases"])
        put(self.conv2_gamma, p["conv2.gamma"]); put(self.conv2_beta, p["conv2.beta"])
        put(self.fc_w, p["softmax_linear.weights"]); put(self.fc_b, p["softmax_linear.biases"])
        for l, lp in enumerate(layers):
            put(self.rnn_W[l], torch.cat([lp["fw"]["W"], lp["bw"]["W"]], 0))
            put(self.rnn_U[l], torch.stack([lp["fw"]["U"], lp["bw"]["U"]], 0))
            if self.cell == "gru":
                put(self.rnn_B[l], torch.stack([lp["fw"]["BW"], lp["bw"]["BW"]], 0))
                put(self.rnn_BU[l], torch.stack([lp["fw"]["BU"], lp["bw"]["BU"]], 0))
            else:
                put(self.rnn_B[l], torch.stack([lp["fw"]["B"], lp["bw"]["B"]], 0))

    def oracle_grads(self):
        """gradients in the oracle's naming, as float64 numpy (for parity tests)."""
        g = lambda t: None if t.grad is None else t.grad.detach().double().cpu().numpy()
        out = {"conv1.weights": g(self.conv1_w), "conv1.biases": g(self.conv1_b), "conv1.gamma": g(self.conv1_gamma),
               "conv1.beta": g(self.conv1_beta), "conv2.weights": g(self.conv2_w), "conv2.biases": g(self.conv2_b),
               "conv2.gamma": g(self.conv2_gamma), "conv2.beta": g(self.conv2_beta),
               "softmax_linear.weights": g(self.fc_w), "softmax_linear.biases": g(self.fc_b)}
        for l in range(self.L):
            out["rnn.%d.W" % l] = g(self.rnn_W[l]); out["rnn.%d.U" % l] = g(self.rnn_U[l]); out["rnn.%d.B" % l] = g(self.rnn_B[l])
            if self.cell == "gru":
                out["rnn.%d.BU" % l] = g(self.rnn_BU[l])
        return out

    # ---- forward pieces
    def frontend(self, feats: torch.Tensor) -> torch.Tensor:
        """feats [B,T,161] f32 -> rnn input [T2,B,75*nf] (bf16 when use_bf16)."""
        B, T, Fq = feats.shape
        x = feats.contiguous().view(B, T, Fq, 1)
        odt = torch.bfloat16 if self.use_bf16 else torch.float32
        if self.use_bf16 and self.nf % 8 == 0:
            # tensor-core route: patches gathered in bf16 (BN1 + clip fused into conv2's gather), tcgen05 GEMMs
            # (conv1: patch rows built in shared memory; both convolutions hand the batch-norm sums of their output on)
            if ops.conv1_fused_takes(x, self.conv1_w, 2, 2):
                y1, s1 = ops.conv1_fused(x, self.conv1_w, self.conv1_b, 2, 2)
            else:
                y1, s1 = ops.conv2d_gemm(x, self.conv1_w, self.conv1_b, 2, 2), None
            y2, s2 = ops.conv2d_gemm(y1, self.conv2_w, self.conv2_b, 2, 1, self.conv1_gamma, self.conv1_beta, self.eps, ops.RELU_CAP,
                                     sums_in=s1, want_sums=True)
            return ops.bn_clip_to_rnn_layout(y2, self.conv2_gamma, self.conv2_beta, self.eps, ops.RELU_CAP,
                                             "cf" if self.layout == "NCHW" else "fc", odt, sums=s2)
        else:
            y1 = ops.conv2d_nhwc(x, self.conv1_w, self.conv1_b, 2, 2)
            a1 = ops.bn_clip_nhwc(y1, self.conv1_gamma, self.conv1_beta, self.eps, ops.RELU_CAP)
            y2 = ops.conv2d_nhwc(a1, self.conv2_w, self.conv2_b, 2, 1)
        return ops.bn_clip_to_rnn_layout(y2, self.conv2_gamma, self.conv2_beta, self.eps, ops.RELU_CAP,
                                         "cf" if self.layout == "NCHW" else "fc", odt)

    def _layer(self, x, l, dirs, reverse_mask, out_dtype):
        H, G = self.H, (3 if self.cell == "gru" else 1)
        if dirs == (0, 1):
            W, U, Bv = self.rnn_W[l], self.rnn_U[l], self.rnn_B[l]
            BU = self.rnn_BU[l] if self.cell == "gru" else None
        else:
            d = dirs[0]
            W, U, Bv = self.rnn_W[l][d * G * H:(d + 1) * G * H], self.rnn_U[l][d:d + 1], self.rnn_B[l][d:d + 1]
            BU = self.rnn_BU[l][d:d + 1] if self.cell == "gru" else None
        if self.cell == "gru":
            return ops.bigru_layer(x, W, U, Bv, BU, self._lens, reverse_mask, self.use_bf16, out_dtype)
        return ops.birnn_layer(x, W, U, Bv, self._lens, reverse_mask, self.use_bf16, out_dtype)

    @torch.no_grad()
    def _layer_eval(self, x, l, rnn_lens_dev):
        """One bi-directional layer of the 'stacked' wiring on its own (the per-layer error report of the parity tests)."""
        self._lens = rnn_lens_dev
        return self._layer(x, l, (0, 1), 0b10, torch.bfloat16 if self.use_bf16 else torch.float32)

    def recurrent_stack(self, x: torch.Tensor, rnn_lens_dev: torch.Tensor) -> torch.Tensor:
        self._lens = rnn_lens_dev
        act = torch.bfloat16 if self.use_bf16 else torch.float32
        if self.blob_engine:
            from .mkldnn_rnn_op import _MkldnnRNNRelu
            layers = [self.L - 1] if self.wiring == "as_written" else range(self.L)
            for l in layers:
                op = _MkldnnRNNRelu(1, self.H, x.shape[2])
                fw = op.sequence(x, self._lens, self.rnn_blob[2 * l], False, self.use_bf16)
                bw = op.sequence(x, self._lens, self.rnn_blob[2 * l + 1], True, self.use_bf16)
                x = _Add.apply(fw, bw)
            return x
        if self.wiring == "as_written":                      # only the last layer is live
            return self._layer(x, self.L - 1, (0, 1), 0b10, act)
        if self.wiring == "stacked":
            B = x.shape[1]
            if self.use_bf16 and ops.rec_stack_supported(self.cell, B, self.H, 2):
                # bf16 tensor-core path: the whole stack as one autograd node (backward overlaps dW/dU with recurrences)
                if self.cell == "gru":
                    lp = [(self.rnn_W[l], self.rnn_U[l], self.rnn_B[l], self.rnn_BU[l]) for l in range(self.L)]
                else:
                    lp = [(self.rnn_W[l], self.rnn_U[l], self.rnn_B[l]) for l in range(self.L)]
                return ops.rec_stack(x, self._lens, self.cell, 0b10, act, lp)
            for l in range(self.L):
                x = self._layer(x, l, (0, 1), 0b10, act)
            return x
        fw = x                                               # two_stack (src/deepSpeech.py:167-185)
        for l in range(self.L):
            fw = self._layer(fw, l, (0,), 0b0, act)
        if not self.bidirectional:
            return fw
        bw = x
        for l in range(self.L):
            bw = self._layer(bw, l, (1,), 0b1, act)
        return _Add.apply(fw, bw)

    def forward(self, feats: torch.Tensor, seq_lens) -> torch.Tensor:
        """-> logits [T2,B,29] fp32."""
        x = self.frontend(feats)
        T2, B, _ = x.shape
        lens = torch.from_numpy(np.minimum(get_rnn_seqlen(seq_lens), T2).astype(np.int32)).to(x.device, non_blocking=True)
        h = self.recurrent_stack(x, lens)
        logits = ops.linear(h.reshape(T2 * B, self.H), self.fc_w, self.fc_b, self.use_bf16)
        return logits.view(T2, B, NUM_CLASSES)


class _Add(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        return ops.add2(a.contiguous(), b.contiguous(), a.dtype)

    @staticmethod
    def backward(ctx, g):
        return g, g
