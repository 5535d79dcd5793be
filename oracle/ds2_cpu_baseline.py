"""fp32 PyTorch-CPU restatement of one reference training step, for TIMING only (bench.py's
`cpu_baseline` and `--impl reference` legs).  TEST/BENCH INFRASTRUCTURE - never imported by the product.

The reference's own TensorFlow-1.x CPU path cannot run in this image (no tensorflow, py2-only source;
BASELINE.md section 2), so this file follows the same call sites with torch-CPU (MKL + oneDNN - the
descendants of the reference's `mkl` engine): src/deepSpeech_NCHW.py:110-201 (conv + batch_norm2 + clip,
NCHW transposes, stacked bi-RNN, FC), src/custom_ops.py:48-72 (per-timestep cell: two matmuls + seq-BN
scale + clipped ReLU inside a python while-loop like tf.nn.bidirectional_dynamic_rnn), :95 (fw+bw),
src/deepSpeech_NCHW.py:225-226 (CTC, batch mean), src/deepSpeech_train.py:287 (autograd backward).
kind = "port".  PARITY UNPINNED (see oracle/ds2_oracle.py header)."""
from __future__ import annotations

import math
import time

import numpy as np
import torch
import torch.nn.functional as F

from . import ds2_oracle as O


def make_params(num_filters=32, num_hidden=1760, num_layers=7, seed=1234):
    p, layers = O.init_params(np.random.default_rng(seed), num_filters, num_hidden, num_layers, dtype=np.float32)
    for t in p.values():
        t.requires_grad_(True)
    for lp in layers:
        for d in lp.values():
            for t in d.values():
                t.requires_grad_(True)
    return p, layers


def _bn_clip(x, gamma, beta, eps):
    y = F.batch_norm(x, None, None, gamma, beta, training=True, eps=eps)
    return torch.clamp(y, 0.0, O.RELU_CAP)


def _direction(x, cp, reverse):
    T, B, _ = x.shape
    H = cp["U"].shape[0]
    s = 1.0 / math.sqrt(1.0 + O.SBN_EPS)
    h = x.new_zeros(B, H)
    outs = [None] * T
    Wt, Ut = cp["W"].t(), cp["U"].t()
    for t in (range(T - 1, -1, -1) if reverse else range(T)):
        resi = torch.mm(x[t], Wt) * s                      # custom_ops.py:61-63 (one matmul per timestep, as TF does)
        h = torch.clamp(torch.addmm(resi + cp["B"], h, Ut), 0.0, O.RELU_CAP)
        outs[t] = h
    return torch.stack(outs, 0)


def train_step(feats, sparse, seq_lens, p, layers):
    """fwd + bwd + CTC of one dummy batch (uniform lengths, as the dummy generator produces)."""
    x = torch.from_numpy(feats).unsqueeze(1)
    y = F.conv2d(x, p["conv1.weights"].permute(3, 2, 0, 1), p["conv1.biases"], stride=(2, 2))
    y = _bn_clip(y, p["conv1.gamma"], p["conv1.beta"], 1e-3)
    y = F.conv2d(y, p["conv2.weights"].permute(3, 2, 0, 1), p["conv2.biases"], stride=(2, 1))
    y = _bn_clip(y, p["conv2.gamma"], p["conv2.beta"], 1e-3)
    B, C, T2, F2 = y.shape
    h = y.permute(2, 0, 1, 3).reshape(T2, B, C * F2)
    for lp in layers:
        h = _direction(h, lp["fw"], False) + _direction(h, lp["bw"], True)
    logits = (h.reshape(-1, h.shape[2]) @ p["softmax_linear.weights"].t() + p["softmax_linear.biases"]).reshape(T2, B, -1)
    seqs = O.sparse_to_label_seqs(*sparse)
    tl = torch.tensor([len(s) for s in seqs])
    loss = F.ctc_loss(torch.log_softmax(logits.float(), 2), torch.tensor(sum(seqs, []), dtype=torch.long),
                      torch.as_tensor(O.get_rnn_seqlen(seq_lens), dtype=torch.long), tl, blank=O.NUM_CLASSES - 1,
                      reduction="none").mean()
    leaves = list(p.values()) + [t for lp in layers for d in lp.values() for t in d.values()]
    torch.autograd.grad(loss, leaves, allow_unused=True)
    return float(loss.detach())


class StepTimer(object):
    """Parameters are built ONCE (89 M values for the train.sh model: seconds of work that is not part of a step);
    `run` then times training steps on one dummy bucket the way the reference does (src/deepSpeech_train.py:306-332:
    wall time of the step including host batch generation, the first `warmup` steps discarded)."""

    def __init__(self, threads, num_hidden=1760, num_layers=7, num_filters=32, seed=1234):
        torch.set_num_threads(threads)
        self.threads = threads
        self.p, self.layers = make_params(num_filters, num_hidden, num_layers, seed)
        self.rng = np.random.default_rng(seed)

    def step(self, batch_size, bucket):
        t0 = time.perf_counter()
        feats, idx, vals, shape, lens = O.dummy_batch(self.rng, batch_size, bucket)
        loss = train_step(feats, (idx, vals, shape), lens, self.p, self.layers)
        return time.perf_counter() - t0, loss

    def run(self, batch_size, bucket, steps, warmup):
        """-> dict(value = audio-s/s from the MEAN step time (the reference's formula), ms_per_step, min/median/max ms, loss)"""
        times, loss = [], float("nan")
        for i in range(warmup + steps):
            dt, loss = self.step(batch_size, bucket)
            if i >= warmup:
                times.append(dt)
        t = np.asarray(times)
        audio_s = batch_size * O.UTT_LENGTHS[bucket] / 100.0
        return {"value": audio_s / float(t.mean()), "ms_per_step": float(t.mean()) * 1e3, "ms_min": float(t.min()) * 1e3,
                "ms_median": float(np.median(t)) * 1e3, "ms_max": float(t.max()) * 1e3, "loss": loss, "audio_s_per_step": audio_s}


def time_steps(batch_size, bucket, steps, warmup, threads, num_hidden=1760, num_layers=7, num_filters=32, seed=1234):
    """-> (audio_seconds_per_second, ms_per_step, loss).  Step time includes host batch generation, as the
    reference's own timing does (src/deepSpeech_train.py:306-324)."""
    r = StepTimer(threads, num_hidden, num_layers, num_filters, seed).run(batch_size, bucket, steps, warmup)
    return r["value"], r["ms_per_step"], r["loss"]


def time_ctc(T, B, L, steps, threads, seed=7):
    """fp32 F.ctc_loss fwd+bwd on the config-2 sweep shapes -> algorithmic GB/s (2*T*B*K*4 bytes)."""
    torch.set_num_threads(threads)
    rng = np.random.default_rng(seed)
    x = torch.tensor(rng.standard_normal((T, B, 29)), dtype=torch.float32, requires_grad=True)
    lab = torch.tensor(rng.integers(1, 28, size=(B, L)), dtype=torch.long)
    best = 1e30
    for _ in range(steps):
        t0 = time.perf_counter()
        loss = F.ctc_loss(torch.log_softmax(x, 2), lab, torch.full((B,), T, dtype=torch.long), torch.full((B,), L, dtype=torch.long),
                          blank=28, reduction="sum", zero_infinity=True)
        torch.autograd.grad(loss, x)
        best = min(best, time.perf_counter() - t0)
    return 2.0 * T * B * 29 * 4 / best / 1e9
