"""Optimizer step of the reference's training loop, fused: Adam + staircase exponential learning-rate decay +
exponential moving average of all trainables (src/deepSpeech_train.py:231-250 set_learning_rate, :430
tf.train.AdamOptimizer(learning_rate), :457-465 apply_gradients / ExponentialMovingAverage / tf.group).
One kernel launch per step for all parameters (ds2_adam_ema_step).  SURVEY 8(f) rank 1: not part of the timed
metric (fwd + bwd + CTC), it is the step that follows it."""
from __future__ import annotations

import torch

from . import ops
from ._lib import check, lib

NUM_PER_EPOCH_FOR_TRAIN = 28535          # src/deepSpeech_input.py:15, read at src/deepSpeech_train.py:240


def set_learning_rate(initial_lr, batch_size, num_epochs_per_decay, lr_decay_factor, num_per_epoch=NUM_PER_EPOCH_FOR_TRAIN):
    """-> (fn(global_step) -> lr, decay_steps), src/deepSpeech_train.py:231-250."""
    num_batches_per_epoch = num_per_epoch / batch_size
    decay_steps = max(1, int(num_batches_per_epoch * num_epochs_per_decay))

    def lr(global_step: int) -> float:
        return float(initial_lr) * float(lr_decay_factor) ** (int(global_step) // decay_steps)
    return lr, decay_steps


class AdamEMA:
    """tf.train.AdamOptimizer(lr) + tf.train.ExponentialMovingAverage(moving_avg_decay, global_step) over `params`
    (CUDA fp32 tensors with .grad).  `lr` is a float or a function of the global step (see set_learning_rate)."""

    def __init__(self, params, lr, moving_avg_decay=0.9999, beta1=0.9, beta2=0.999, epsilon=1e-8, ema=True):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("AdamEMA: no trainable parameters")
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("AdamEMA needs contiguous fp32 CUDA parameters (no CPU fallback)")
        self.lr, self.decay = lr, float(moving_avg_decay)
        self.beta1, self.beta2, self.eps = beta1, beta2, epsilon
        self.global_step = 0
        self.m = [torch.zeros_like(p) for p in self.params]
        self.v = [torch.zeros_like(p) for p in self.params]
        self.shadow = [p.detach().clone() for p in self.params] if ema else None     # EMA starts at the initial value
        self._table = None
        self._grad_ptrs = None
        self._chunk = lib.ds2_adam_ema_chunk()

    def _build_table(self):
        rows, c0 = [], 0
        for i, p in enumerate(self.params):
            if p.grad is None:
                raise RuntimeError("AdamEMA.step: parameter %d has no gradient" % i)
            g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
            if g is not p.grad:
                p.grad = g
            n = p.numel()
            rows.append([p.data_ptr(), g.data_ptr(), self.m[i].data_ptr(), self.v[i].data_ptr(),
                         self.shadow[i].data_ptr() if self.shadow is not None else 0, n, c0])
            c0 += (n + self._chunk - 1) // self._chunk
        self._grad_ptrs = [r[1] for r in rows]
        self._n_chunks = c0
        self._table = torch.tensor(rows, dtype=torch.int64).to(self.params[0].device)

    def learning_rate(self) -> float:
        return float(self.lr(self.global_step)) if callable(self.lr) else float(self.lr)

    def step(self, grad_scale: float = 1.0):
        if self._table is None or self._grad_ptrs != [p.grad.data_ptr() if p.grad is not None else 0 for p in self.params]:
            self._build_table()                  # gradients were re-allocated (e.g. zero_grad(set_to_none=True))
        lr = self.learning_rate()                # the rate of this step is read before global_step is incremented
        self.global_step += 1
        n = self.global_step
        keep = min(self.decay, (1.0 + n) / (10.0 + n))
        check(lib.ds2_adam_ema_step(self._table.data_ptr(), len(self.params), self._n_chunks, lr, self.beta1, self.beta2,
                                    self.eps, n, keep, float(grad_scale), torch.cuda.current_stream().cuda_stream),
              "ds2_adam_ema_step")
        ops.invalidate_weight_cache()          # the parameters changed behind autograd's back: bf16 copies are stale

    def zero_grad(self, set_to_none: bool = True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()
