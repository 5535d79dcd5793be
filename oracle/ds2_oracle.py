"""CPU oracle for the DeepSpeech2 training hot path (conv front-end -> stacked bi-RNN -> FC -> CTC).

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import it.  The product
(deepspeech_b200/) never imports anything under oracle/ and has no CPU fallback.

PARITY UNPINNED: the reference (yxlao/deepSpeech) ships no tests, no golden vectors and no
fixtures, and every FLOP of this path executes inside TensorFlow 1.1/1.2
(README.md:10,31), an un-vendored third-party dependency that is absent from /root/reference
and from this image (python2-only source, no `tensorflow` module).  This file therefore
restates, in float64 torch-CPU / numpy, the reference's *call sites* line by line plus the
published semantics of the TF ops they call (conv2d VALID cross-correlation with HWIO
filters; fused_batch_norm with biased batch variance; dynamic_rnn zero-output/frozen-state
past seq_len; reverse_sequence over the valid prefix for the backward direction; ctc_loss
with blank = num_classes-1, softmax inside, unnormalised per-utterance NLL).  It is pinned
against (i) closed-form known answers and (ii) an independent implementation
(torch.nn.functional.ctc_loss / conv2d / nn.RNN / nn.GRU, CPU fp64) in tests/test_oracle.py.

All citations `file:line` are relative to /root/reference/.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

NUM_CLASSES = 29          # src/deepSpeech_input.py:13-14 (28 characters + blank)
FREQ_BINS = 161           # src/deepSpeech_dummy.py:13
RELU_CAP = 20.0           # src/custom_ops.py:71 ; src/deepSpeech_NCHW.py:133,157
SBN_EPS = 1e-5            # src/custom_ops.py:198
DT = torch.float64


# ----------------------------------------------------------------------------- lengths
def get_rnn_seqlen(seq_lens):
    """src/deepSpeech.py:38-48 : ceil((ceil((len-19)/2) - 9)/2) evaluated in float64."""
    s = np.asarray(seq_lens, dtype=np.float64)
    s = np.ceil((s - 19.0) / 2.0)
    s = np.ceil((s - 9.0) / 2.0)
    return s.astype(np.int32)


def conv_out_lens(T: int) -> Tuple[int, int]:
    """VALID conv output sizes along time: conv1 20-tap stride 2, conv2 10-tap stride 2
    (src/deepSpeech_NCHW.py:120-123, :142-146)."""
    t1 = (T - 20) // 2 + 1
    t2 = (t1 - 10) // 2 + 1
    return t1, t2


# ----------------------------------------------------------------------------- elementwise
def relux(x: torch.Tensor, capping=None) -> torch.Tensor:
    """src/custom_ops.py:99-104.  capping=None raises NameError in the reference (quirk Q5);
    the oracle (like the product) treats it as plain ReLU."""
    y = torch.relu(x)
    if capping is not None:
        y = torch.clamp(y, max=float(capping))
    return y


def batch_norm_train(x: torch.Tensor, gamma, beta, eps: float, channel_axis: int) -> torch.Tensor:
    """Training-mode batch norm with *biased* batch variance over every axis but the channel
    axis.  Restates both src/custom_ops.py:107-160 (`batch_norm2`: tf.nn.fused_batch_norm,
    eps 1e-3, is_training=True) and src/custom_ops.py:163-181 (`batch_norm`: tf.nn.moments over
    [0,1,2] + tf.nn.batch_normalization, eps 1e-5)."""
    axes = [a for a in range(x.dim()) if a != channel_axis % x.dim()]
    mean = x.mean(dim=axes, keepdim=True)
    var = ((x - mean) ** 2).mean(dim=axes, keepdim=True)
    shape = [1] * x.dim()
    shape[channel_axis] = -1
    return (x - mean) / torch.sqrt(var + eps) * gamma.reshape(shape) + beta.reshape(shape)


def seq_batch_norm_as_written(x: torch.Tensor) -> torch.Tensor:
    """src/custom_ops.py:184-199 as written (quirk Q2): the moving statistics are never
    updated (the assign ops at :196-197 are dropped), so the op normalises with
    moving_mean = 0, moving_variance = 1, beta = 0, gamma = 1: x / sqrt(1 + 1e-5)."""
    return x / math.sqrt(1.0 + SBN_EPS)


# ----------------------------------------------------------------------------- front-end
def conv_frontend(feats: torch.Tensor, p: Dict[str, torch.Tensor], layout: str = "NCHW",
                  return_intermediates: bool = False):
    """feats [B,T,161] -> rnn_input [T2,B,75*nf].

    NCHW: src/deepSpeech_NCHW.py:110-168 (batch_norm2 eps 1e-3, flatten feature = c*75+f).
    NHWC: src/deepSpeech.py:110-164      (batch_norm  eps 1e-5, flatten feature = f*nf+c).
    Filters are HWIO ([20,5,1,nf] / [10,5,nf,nf]); tf.nn.conv2d is a VALID cross-correlation.
    """
    eps = 1e-3 if layout == "NCHW" else 1e-5
    x = feats.to(DT).unsqueeze(1)                              # [B,1,T,F]
    w1 = p["conv1.weights"].to(DT).permute(3, 2, 0, 1)         # HWIO -> OIHW
    y1 = F.conv2d(x, w1, p["conv1.biases"].to(DT), stride=(2, 2))
    a1 = relux(batch_norm_train(y1, p["conv1.gamma"].to(DT), p["conv1.beta"].to(DT), eps, 1), RELU_CAP)
    w2 = p["conv2.weights"].to(DT).permute(3, 2, 0, 1)
    y2 = F.conv2d(a1, w2, p["conv2.biases"].to(DT), stride=(2, 1))
    a2 = relux(batch_norm_train(y2, p["conv2.gamma"].to(DT), p["conv2.beta"].to(DT), eps, 1), RELU_CAP)
    B, C, T2, F2 = a2.shape
    if layout == "NCHW":
        out = a2.permute(2, 0, 1, 3).reshape(T2, B, C * F2)    # deepSpeech_NCHW.py:166-168
    else:
        out = a2.permute(2, 0, 3, 1).reshape(T2, B, F2 * C)    # deepSpeech.py:162-164
    if return_intermediates:
        return out, dict(y1=y1, a1=a1, y2=y2, a2=a2)
    return out


# ----------------------------------------------------------------------------- recurrent cells
def rnn_cell_preact(x_t, h_prev, W, U, b, sbn: str = "as_written"):
    """src/custom_ops.py:59-68 : the argument of the clipped ReLU, sbn(x W^T) + h U^T + b."""
    resi = x_t @ W.t()
    if sbn == "as_written":
        resi = seq_batch_norm_as_written(resi)
    resu = h_prev @ U.t()
    return resi + resu + b


def rnn_cell_step(x_t, h_prev, W, U, b, sbn: str = "as_written"):
    """src/custom_ops.py:48-72 : h = min(relu(sbn(x W^T) + h U^T + b), 20)."""
    return relux(rnn_cell_preact(x_t, h_prev, W, U, b, sbn), RELU_CAP)


def gru_cell_step(x_t, h_prev, W, U, bw, bu, sbn: str = "as_written"):
    """GRU is ABSENT from the reference; build-defined as the cuDNN/torch.nn.GRU
    gate form (gate order r,z,n) with the reference's seq-BN applied to W x:
      r,z = sigmoid(sbn(Wx)+bw + Uh+bu) ; n = tanh(sbn(W_n x)+bw_n + r*(U_n h + bu_n)) ;
      h' = (1-z)*n + z*h."""
    H = h_prev.shape[-1]
    gi = x_t @ W.t()
    if sbn == "as_written":
        gi = seq_batch_norm_as_written(gi)
    gi = gi + bw
    gh = h_prev @ U.t() + bu
    r = torch.sigmoid(gi[:, :H] + gh[:, :H])
    z = torch.sigmoid(gi[:, H:2 * H] + gh[:, H:2 * H])
    n = torch.tanh(gi[:, 2 * H:] + r * gh[:, 2 * H:])
    return (1.0 - z) * n + z * h_prev


def rnn_direction(x, lens, cell_params, reverse: bool, cell: str = "relu_clip", sbn: str = "as_written", preacts=None):
    """One direction of tf.nn.bidirectional_dynamic_rnn (src/custom_ops.py:91-93,
    src/deepSpeech.py:179-183), time-major, zero initial state (custom_ops.py:89-90).
    dynamic_rnn semantics (TF-internal): for t >= len[b] the emitted output is zero and the
    state is copied through; the backward direction runs over reverse_sequence(x, len), i.e.
    over the valid prefix only, and its outputs are reversed back.
    preacts (relu_clip only): a list of length T that receives the pre-activation tensor of every step (kept in
    the autograd graph with retain_grad, so that d loss / d pre-activation can be read after backward())."""
    T, B, _ = x.shape
    H = cell_params["U"].shape[1]
    lens_t = torch.as_tensor(np.asarray(lens), dtype=torch.int64)
    h = torch.zeros(B, H, dtype=x.dtype)
    outs = [None] * T
    order = range(T - 1, -1, -1) if reverse else range(T)
    for t in order:
        if cell == "relu_clip" and preacts is not None:
            z = rnn_cell_preact(x[t], h, cell_params["W"], cell_params["U"], cell_params["B"], sbn)
            if z.requires_grad:
                z.retain_grad()
            preacts[t] = z
            hn = relux(z, RELU_CAP)
        elif cell == "relu_clip":
            hn = rnn_cell_step(x[t], h, cell_params["W"], cell_params["U"], cell_params["B"], sbn)
        elif cell == "relu_plain":
            # src/mkldnn_rnn_op.py:22-45 (MkldnnRNNRelu: plain ReLU, no seq-BN, no clip; build-defined)
            hn = torch.relu(x[t] @ cell_params["W"].t() + h @ cell_params["U"].t() + cell_params["B"])
        else:
            hn = gru_cell_step(x[t], h, cell_params["W"], cell_params["U"], cell_params["BW"],
                               cell_params["BU"], sbn)
        valid = (t < lens_t).to(x.dtype).unsqueeze(1)
        # backward direction: state stays at its zero init until the first valid frame
        h = valid * hn + (1.0 - valid) * (torch.zeros_like(h) if reverse else h)
        outs[t] = valid * hn
    return torch.stack(outs, 0)


def stacked_brnn(x, lens, layers: Sequence[Dict[str, Dict[str, torch.Tensor]]], wiring: str = "stacked",
                 cell: str = "relu_clip", sbn: str = "as_written", bidirectional: bool = True):
    """Three topologies:
      'as_written' : src/custom_ops.py:86-96 verbatim - every layer reads the ORIGINAL inputs
                     and only the last layer's fw+bw reaches the output.
      'stacked'    : the docstring's intent (custom_ops.py:76-85): layer i+1 reads fw_i+bw_i.
      'two_stack'  : src/deepSpeech.py:167-185 - MultiRNNCell: two independent L-deep
                     uni-directional stacks (one over reversed time) summed once at the end;
                     bidirectional=False gives the 'uni-dir' branch (:173-177).
    """
    if wiring == "as_written":
        lp = layers[-1]
        return rnn_direction(x, lens, lp["fw"], False, cell, sbn) + rnn_direction(x, lens, lp["bw"], True, cell, sbn)
    if wiring == "stacked":
        cur = x
        for lp in layers:
            cur = rnn_direction(cur, lens, lp["fw"], False, cell, sbn) + rnn_direction(cur, lens, lp["bw"], True, cell, sbn)
        return cur
    if wiring == "two_stack":
        fw = x
        for lp in layers:
            fw = rnn_direction(fw, lens, lp["fw"], False, cell, sbn)
        if not bidirectional:
            return fw
        bw = x
        for lp in layers:
            bw = rnn_direction(bw, lens, lp["bw"], True, cell, sbn)
        return fw + bw
    raise ValueError(wiring)


def fc_logits(h, W, b):
    """src/deepSpeech_NCHW.py:188-199 : reshape [-1,H] @ W^T[29,H] + b -> [T,B,29]."""
    T, B, H = h.shape
    return (h.reshape(-1, H) @ W.t() + b).reshape(T, B, -1)


# ----------------------------------------------------------------------------- CTC
def sparse_to_label_seqs(indices, values, dense_shape, num_classes: int = NUM_CLASSES,
                         preprocess_collapse_repeated: bool = True) -> List[List[int]]:
    """Label pre-processing of tf.nn.ctc_loss as called at src/deepSpeech_NCHW.py:225 with the
    SparseTensor built at src/deepSpeech_train.py:437 from src/deepSpeech_dummy.py:89-98
    (zeros were already dropped by _dense_to_sparse).  Per utterance, in index order:
    adjacent repeats are collapsed when preprocess_collapse_repeated; a label >= num_classes-1
    ends the sequence and a valid label after it is an error (TF-internal semantics, Q8)."""
    B = int(dense_shape[0])
    seqs: List[List[int]] = [[] for _ in range(B)]
    raw: List[List[int]] = [[] for _ in range(B)]
    for (b, _c), v in zip(np.asarray(indices).reshape(-1, 2), np.asarray(values).reshape(-1)):
        raw[int(b)].append(int(v))
    for b in range(B):
        finished = False
        lab = raw[b]
        for j, v in enumerate(lab):
            if j == 0 or not preprocess_collapse_repeated or v != lab[j - 1]:
                if v >= num_classes - 1 or v < 0:
                    finished = True
                else:
                    if finished:
                        raise ValueError("Saw a non-null label (index >= num_classes - 1) following a "
                                         "null label, batch: %d" % b)
                    seqs[b].append(v)
    return seqs


def _lprime(label: Sequence[int], blank: int) -> np.ndarray:
    lp = np.full(2 * len(label) + 1, blank, dtype=np.int64)
    lp[1::2] = np.asarray(label, dtype=np.int64)
    return lp


def ctc_loss_and_grad_np(logits: np.ndarray, labels: List[List[int]], seq_lens, blank: int = NUM_CLASSES - 1):
    """float64 log-space alpha/beta CTC, TF conventions (TF-internal ctc_loss_calculator):
    alpha_t(s) includes y_t(l'_s); beta_t(s) excludes it, so sum_s alpha_t(s)beta_t(s) = p for
    every t; loss_b = -log p; d loss_b / d logit[t,b,k] = y[t,k] - (1/p) sum_{s:l'_s=k} alpha_t(s)beta_t(s),
    zero for t >= seq_len[b].  Raises like TF if the utterance is too short for its label."""
    logits = np.asarray(logits, dtype=np.float64)
    T, B, K = logits.shape
    m = logits.max(axis=2, keepdims=True)
    logy = logits - m - np.log(np.exp(logits - m).sum(axis=2, keepdims=True))
    loss = np.zeros(B)
    grad = np.zeros_like(logits)
    NEG = -np.inf
    for b in range(B):
        Tb = int(seq_lens[b])
        lab = labels[b]
        L = len(lab)
        rep = sum(1 for i in range(1, L) if lab[i] == lab[i - 1])
        if Tb < L + rep:
            raise ValueError("Not enough time for target transition sequence (required: %d, available: %d) "
                             "batch: %d" % (L + rep, Tb, b))
        if Tb == 0:
            if L > 0:
                raise ValueError("Not enough time")
            continue
        lp = _lprime(lab, blank)
        S = lp.shape[0]
        ly = logy[:Tb, b, :][:, lp]                      # [Tb,S]
        can_skip = np.zeros(S, dtype=bool)
        can_skip[2:] = (lp[2:] != blank) & (lp[2:] != lp[:-2])
        alpha = np.full((Tb, S), NEG)
        alpha[0, 0] = ly[0, 0]
        if S > 1:
            alpha[0, 1] = ly[0, 1]
        for t in range(1, Tb):
            a = alpha[t - 1]
            a1 = np.concatenate(([NEG], a[:-1]))[:S]
            a2 = np.where(can_skip, np.concatenate(([NEG, NEG], a[:-2]))[:S], NEG)
            alpha[t] = np.logaddexp(np.logaddexp(a, a1), a2) + ly[t]
        beta = np.full((Tb, S), NEG)
        beta[Tb - 1, S - 1] = 0.0
        if S > 1:
            beta[Tb - 1, S - 2] = 0.0
        skip_to = np.zeros(S, dtype=bool)                # s -> s+2 allowed
        skip_to[:-2] = can_skip[2:]
        for t in range(Tb - 2, -1, -1):
            bn = beta[t + 1] + ly[t + 1]
            b1 = np.concatenate((bn[1:], [NEG]))[:S]
            b2 = np.where(skip_to, np.concatenate((bn[2:], [NEG, NEG]))[:S], NEG)
            beta[t] = np.logaddexp(np.logaddexp(bn, b1), b2)
        logp = np.logaddexp(alpha[Tb - 1, S - 1], alpha[Tb - 1, S - 2]) if S > 1 else alpha[Tb - 1, 0]
        loss[b] = -logp
        ab = np.exp(alpha + beta - logp)                 # [Tb,S]
        occ = np.zeros((Tb, K))
        np.add.at(occ, (np.arange(Tb)[:, None].repeat(S, 1), np.broadcast_to(lp, (Tb, S))), ab)
        grad[:Tb, b, :] = np.exp(logy[:Tb, b, :]) - occ
    return loss, grad


def ctc_loss_torch(logits: torch.Tensor, labels: List[List[int]], seq_lens, blank: int = NUM_CLASSES - 1):
    """Differentiable (autograd) float64 restatement of the same recursion; used so the whole
    model oracle can be back-propagated (src/deepSpeech_train.py:287).  Returns loss[B]."""
    T, B, K = logits.shape
    logy = torch.log_softmax(logits, dim=2)
    NEG = torch.tensor(-1e300, dtype=logits.dtype)
    losses = []
    for b in range(B):
        Tb = int(seq_lens[b])
        lab = labels[b]
        lp = torch.as_tensor(_lprime(lab, blank))
        S = lp.shape[0]
        ly = logy[:Tb, b, :][:, lp]
        can_skip = torch.zeros(S, dtype=torch.bool)
        if S > 2:
            can_skip[2:] = (lp[2:] != blank) & (lp[2:] != lp[:-2])
        a = torch.full((S,), -1e300, dtype=logits.dtype)
        init = [ly[0, 0]] + ([ly[0, 1]] if S > 1 else [])
        a = torch.cat([torch.stack(init), a[len(init):]])
        for t in range(1, Tb):
            a1 = torch.cat([NEG.reshape(1), a[:-1]])
            a2 = torch.where(can_skip, torch.cat([NEG.reshape(1), NEG.reshape(1), a[:-2]])[:S], NEG)
            a = torch.logsumexp(torch.stack([a, a1, a2]), dim=0) + ly[t]
        logp = torch.logsumexp(a[-2:], dim=0) if S > 1 else a[0]
        losses.append(-logp)
    return torch.stack(losses)


# ----------------------------------------------------------------------------- whole model
def model_forward(feats, seq_lens, p: Dict[str, torch.Tensor], layers, cfg) -> torch.Tensor:
    """inference(): src/deepSpeech_NCHW.py:81-201 (layout NCHW) / src/deepSpeech.py:81-203 (NHWC)."""
    x = conv_frontend(feats, p, cfg.get("layout", "NCHW"))
    lens = get_rnn_seqlen(seq_lens)
    h = stacked_brnn(x, lens, layers, cfg.get("wiring", "stacked"), cfg.get("cell", "relu_clip"),
                     cfg.get("sbn", "as_written"), cfg.get("bidirectional", True))
    return fc_logits(h, p["softmax_linear.weights"].to(DT), p["softmax_linear.biases"].to(DT))


def model_loss(feats, seq_lens, label_seqs, p, layers, cfg) -> Tuple[torch.Tensor, torch.Tensor]:
    """loss(): src/deepSpeech_NCHW.py:204-228 : batch mean of per-utterance CTC NLL (Q13)."""
    logits = model_forward(feats, seq_lens, p, layers, cfg)
    per_utt = ctc_loss_torch(logits, label_seqs, get_rnn_seqlen(seq_lens))
    return per_utt.mean(), logits


# ----------------------------------------------------------------------------- weights & dummy data
def init_params(rng: np.random.Generator, num_filters=32, num_hidden=1760, num_layers=7, cell="relu_clip",
                wiring="stacked", dtype=np.float64):
    """Initialisers of src/helper_routines.py:31-78 and the call sites: conv/FC weights =
    variance_scaling(factor 2, FAN_IN, truncated normal) (deepSpeech_NCHW.py:112-115,138-141,
    185-187), conv biases -0.05 (:124-126,:147-150), FC bias 0 (:191-193), BN gamma 1 / beta 0
    (custom_ops.py:148-153), RNN W/U = tf.get_variable default (glorot uniform, TF-internal),
    RNN bias 0 (custom_ops.py:68-70)."""
    def trunc_normal(shape, std):
        out = rng.standard_normal(shape)
        bad = np.abs(out) > 2.0
        while bad.any():
            out[bad] = rng.standard_normal(int(bad.sum()))
            bad = np.abs(out) > 2.0
        return out * std

    def vs(shape):          # fan_in = prod(shape[:-1]) for conv, shape[-2] for 2-D (TF contrib rule)
        fan_in = float(np.prod(shape[:-1])) if len(shape) == 4 else float(shape[-2])
        return trunc_normal(shape, math.sqrt(1.3 * 2.0 / fan_in))

    def glorot(shape):
        lim = math.sqrt(6.0 / (shape[0] + shape[1]))
        return rng.uniform(-lim, lim, size=shape)

    nf, H = num_filters, num_hidden
    p = {
        "conv1.weights": vs((20, 5, 1, nf)), "conv1.biases": np.full(nf, -0.05),
        "conv1.gamma": np.ones(nf), "conv1.beta": np.zeros(nf),
        "conv2.weights": vs((10, 5, nf, nf)), "conv2.biases": np.full(nf, -0.05),
        "conv2.gamma": np.ones(nf), "conv2.beta": np.zeros(nf),
        "softmax_linear.weights": vs((NUM_CLASSES, H)), "softmax_linear.biases": np.zeros(NUM_CLASSES),
    }
    G = 3 if cell == "gru" else 1
    layers = []
    for l in range(num_layers):
        I = 75 * nf if (l == 0 or wiring == "as_written") else H
        lp = {}
        for d in ("fw", "bw"):
            q = {"W": glorot((G * H, I)), "U": glorot((G * H, H))}
            if cell == "gru":
                q["BW"] = np.zeros(G * H)
                q["BU"] = np.zeros(G * H)
            else:
                q["B"] = np.zeros(H)
            lp[d] = {k: torch.tensor(np.asarray(v, dtype=dtype)) for k, v in q.items()}
        layers.append(lp)
    p = {k: torch.tensor(np.asarray(v, dtype=dtype)) for k, v in p.items()}
    return p, layers


UTT_LENGTHS = [100, 200, 100, 400, 500, 600, 700, 800, 900, 1000, 1100, 1200, 1300, 1400, 1500]  # dummy.py:9
COUNTS = [30, 3, 3, 3, 14, 13, 9, 8, 5, 4, 3, 2, 2, 2, 1]                                          # dummy.py:10
LABEL_LENGTHS = [7, 17, 7, 48, 62, 78, 93, 107, 120, 134, 148, 163, 178, 193, 209]                 # dummy.py:11


def dummy_batch(rng: np.random.Generator, batch_size: int, bucket: int):
    """One batch of src/deepSpeech_dummy.py:54-122 for a given bucket: randn features
    [B,T,161] f32, ONE label sequence shared by the whole batch (values randint(0,27), last =
    28), dense -> sparse dropping zeros (:89-98), seq_lens = T."""
    T = UTT_LENGTHS[bucket]
    L = LABEL_LENGTHS[bucket]
    feats = rng.standard_normal((batch_size, T, FREQ_BINS)).astype(np.float32)
    label = rng.integers(0, NUM_CLASSES - 1, size=L)          # random.randint(0, 27) inclusive
    label[L - 1] = NUM_CLASSES - 1
    idx, vals = [], []
    for b in range(batch_size):
        for c in range(L):
            if label[c] != 0:
                idx.append([b, c])
                vals.append(int(label[c]))
    return feats, np.asarray(idx, dtype=np.int64).reshape(-1, 2), np.asarray(vals, dtype=np.int32), \
        np.asarray([batch_size, L], dtype=np.int64), np.full(batch_size, T, dtype=np.int32)


# ------------------------------------------------------------------------------------------------------------------
# Optimizer step.  TF-internal semantics restated from memory (parity unpinned, like the rest):
# tf.train.exponential_decay(staircase=True) at src/deepSpeech_train.py:231-250, tf.train.AdamOptimizer defaults
# (beta1 0.9, beta2 0.999, epsilon 1e-8) at :430 / :457-458, tf.train.ExponentialMovingAverage(decay, num_updates) at
# :461-462.  The reference groups the Adam update and the averaging without a control dependency (:465), so their
# order is undefined there; the build fixes it as "update, then average" with num_updates = updates done so far
# (the order of the CIFAR-10 multi-GPU tutorial this file derives from).
def decayed_learning_rate(initial_lr, global_step, decay_steps, decay_factor):
    """tf.train.exponential_decay(..., staircase=True): lr * factor ** floor(step / decay_steps)."""
    return float(initial_lr) * float(decay_factor) ** (int(global_step) // int(decay_steps))


def ema_decay(decay, num_updates):
    """tf.train.ExponentialMovingAverage: min(decay, (1 + n) / (10 + n))."""
    return min(float(decay), (1.0 + num_updates) / (10.0 + num_updates))


def adam_ema_step_np(w, g, m, v, shadow, step, lr, decay, beta1=0.9, beta2=0.999, eps=1e-8, grad_scale=1.0):
    """One update of one tensor in float64; step = 1 for the first update.  Returns (w, m, v, shadow)."""
    w, g, m, v = (np.asarray(a, np.float64) for a in (w, g, m, v))
    g = g * grad_scale
    m = beta1 * m + (1.0 - beta1) * g
    v = beta2 * v + (1.0 - beta2) * g * g
    lr_t = lr * np.sqrt(1.0 - beta2 ** step) / (1.0 - beta1 ** step)
    w = w - lr_t * m / (np.sqrt(v) + eps)
    if shadow is not None:
        shadow = np.asarray(shadow, np.float64)
        shadow = shadow - (1.0 - ema_decay(decay, step)) * (shadow - w)
    return w, m, v, shadow


# ------------------------------------------------------------------------------------------------------------------
# Evaluation side: greedy CTC decoding and character error rate, src/deepSpeech_test.py:112-136,
# :209-213.  tf.nn.ctc_greedy_decoder (TF-internal, from memory): per-frame argmax (first maximum), merge_repeated=True
# collapses runs, the blank (K-1) is dropped; Levenshtein.distance is the unit-cost edit distance.
ALPHABET = "ABCDEFGHIJKLMNOPQRSTUVWXYZ' "      # src/deepSpeech_test.py:28


def ctc_greedy_decode_np(logits, seq_lens):
    """logits [T,B,K] -> list of B lists of class ids."""
    logits = np.asarray(logits)
    T, B, K = logits.shape
    out = []
    for b in range(B):
        prev, seq = -1, []
        for t in range(min(int(seq_lens[b]), T)):
            c = int(np.argmax(logits[t, b]))
            if c != prev and c != K - 1:
                seq.append(c)
            prev = c
        out.append(seq)
    return out


def edit_distance_py(a, b):
    """Levenshtein distance (unit costs) of two sequences."""
    la, lb = len(a), len(b)
    row = list(range(lb + 1))
    for i in range(1, la + 1):
        prev_diag, row[0] = row[0], i
        for j in range(1, lb + 1):
            cur = min(row[j] + 1, row[j - 1] + 1, prev_diag + (a[i - 1] != b[j - 1]))
            prev_diag, row[j] = row[j], cur
    return row[lb]


def char_error_rates(label_seqs, pred_seqs):
    """distance(label, pred) / len(label) per utterance, src/deepSpeech_test.py:129-130."""
    return [edit_distance_py(l, p) / len(l) for l, p in zip(label_seqs, pred_seqs)]
