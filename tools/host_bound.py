#!/usr/bin/env python
"""Is the training step bound by the host's enqueue rate?  (needs a GPU)  Times, for the bench's train.sh step: the host time
to ENQUEUE a step (no synchronisation inside) and the GPU time per step (events around a run of steps)."""
import os
import sys
import time
from types import SimpleNamespace

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepspeech_b200 import deepSpeech_NCHW as ds
from deepspeech_b200.deepSpeech_dummy import DummyData

B, T = 32, 1500
params = SimpleNamespace(use_fp16=True, num_filters=32, batch_size=B, num_hidden=1760, num_rnn_layers=7, engine="mkl",
                         rnn_type="bi-dir", wiring="stacked", seed=1)
model = ds.model(params)
data = DummyData(B, seed=3)
feats, idx, vals, shape, lens = data.inputs(14)
f = torch.from_numpy(feats).cuda()
sparse = (idx, vals, shape)


def step():
    model.zero_grad(set_to_none=True)
    logits = ds.inference(None, f, lens, params)
    loss = ds.loss(logits, sparse, lens)
    loss.backward()
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
n = 10
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
host = []
e0.record()
for _ in range(n):
    t0 = time.perf_counter()
    step()
    host.append((time.perf_counter() - t0) * 1e3)
e1.record()
torch.cuda.synchronize()
print("frames per utterance %d; GPU %.2f ms per step; host enqueue per step: min %.2f median %.2f max %.2f ms"
      % (feats.shape[1], e0.elapsed_time(e1) / n, min(host), float(np.median(host)), max(host)))
# one step after a synchronize: the GPU starts with an empty queue
ts = []
for _ in range(5):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    step()
    b.record()
    torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
print("a single step after a synchronize: %.2f ms" % float(np.median(ts)))
