#!/usr/bin/env python
"""Timing of the recurrent stack's backward with and without the side-stream weight gradients (needs a GPU).
usage: tools/stack_overlap.py [cell] [H] [L] [B] [T]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deepspeech_b200 import ops

cell = sys.argv[1] if len(sys.argv) > 1 else "relu_clip"
H = int(sys.argv[2]) if len(sys.argv) > 2 else 1760
L = int(sys.argv[3]) if len(sys.argv) > 3 else 7
B = int(sys.argv[4]) if len(sys.argv) > 4 else 32
T = int(sys.argv[5]) if len(sys.argv) > 5 else 366
dev = "cuda"
G = 3 if cell == "gru" else 1
torch.manual_seed(0)
I0 = 2400
params = []
for l in range(L):
    I = I0 if l == 0 else H
    lp = [torch.randn(2 * G * H, I, device=dev) * (1.0 / I ** 0.5), torch.randn(2, G * H, H, device=dev) * (0.5 / H ** 0.5),
          torch.zeros(2, G * H, device=dev)]
    if cell == "gru":
        lp.append(torch.zeros(2, G * H, device=dev))
    params.append([t.requires_grad_(True) for t in lp])
x = (torch.randn(T, B, I0, device=dev) * 0.5).to(torch.bfloat16).requires_grad_(True)
lens = torch.full((B,), T, dtype=torch.int32, device=dev)
g = torch.randn(T, B, H, device=dev).to(torch.bfloat16)


def run(n=5):
    fw, bw = [], []
    for it in range(n + 2):
        for lp in params:
            for t in lp:
                t.grad = None
        x.grad = None
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        out = ops.rec_stack(x, lens, cell, 0b10, torch.bfloat16, params)
        e[1].record()
        out.backward(g)
        e[2].record()
        torch.cuda.synchronize()
        if it >= 2:
            fw.append(e[0].elapsed_time(e[1]))
            bw.append(e[1].elapsed_time(e[2]))
    return sum(fw) / len(fw), sum(bw) / len(bw)


for ov in (False, True, False, True):
    ops.OVERLAP_WGRAD = ov
    f, b = run()
    print("cell=%s H=%d L=%d B=%d T=%d overlap=%s: forward %.3f ms  backward %.3f ms" % (cell, H, L, B, T, ov, f, b))
# timeline of one backward: start/end of each recurrence and each side batch relative to the first recurrence start
for ov in (False, True):
    ops.OVERLAP_WGRAD = ov
    ops.STACK_TRACE = []
    run(1)
    tr = ops.STACK_TRACE[-2 * L:]
    ops.STACK_TRACE = None
    t0 = tr[0][2]
    print("overlap=%s" % ov)
    for tag, l, a, b in tr:
        print("  %-4s layer %d: start %8.3f ms  end %8.3f ms  (%.3f ms)" % (tag, l, t0.elapsed_time(a), t0.elapsed_time(b), a.elapsed_time(b)))
