#!/usr/bin/env python
"""Standalone CTC loss+grad timing (BASELINE config 2 shapes) through the C ABI; also the ncu target.
usage: tools/ctc_bench.py [T] [L] [B] [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepspeech_b200 import ops
from deepspeech_b200._lib import check, lib

T = int(sys.argv[1]) if len(sys.argv) > 1 else 400
L = int(sys.argv[2]) if len(sys.argv) > 2 else 100
B = int(sys.argv[3]) if len(sys.argv) > 3 else 64
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 20
dev = "cuda"
x = torch.tensor(np.random.default_rng(7).standard_normal((T, B, 29)), dtype=torch.float32, device=dev)
rng = np.random.default_rng(11)
labs = []
for b in range(B):
    r = rng.integers(0, 28, size=L)
    labs.append([int(v) for i, v in enumerate(r) if i == 0 or v != r[i - 1]])
Lm = max(len(r) for r in labs)
lab = np.zeros((B, Lm), np.int32)
for b, r in enumerate(labs):
    lab[b, :len(r)] = r
lab_d = torch.from_numpy(lab).to(dev)
ll_d = torch.tensor([len(r) for r in labs], dtype=torch.int32, device=dev)
sl_d = torch.full((B,), T, dtype=torch.int32, device=dev)
loss, grad = torch.empty(B, device=dev), torch.empty_like(x)
wsb = lib.ds2_ctc_workspace_bytes(T, B, 29, Lm)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
p, st = ops._p, ops._stream


def run():
    check(lib.ds2_ctc_loss_fwd_bwd(p(x), T, B, 29, p(lab_d), Lm, p(ll_d), p(sl_d), None, p(loss), p(grad), p(ws), wsb, st()), "ctc")


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print("CTC T=%d L=%d B=%d: %.1f us  (%.0f ns / lattice step, %.1f GB/s algorithmic)" % (T, L, B, ms * 1e3, ms * 1e6 / T, 2.0 * T * B * 29 * 4 / ms / 1e6))
