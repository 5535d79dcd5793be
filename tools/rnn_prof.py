#!/usr/bin/env python
"""Per-stage latency of the persistent recurrent kernel (needs a GPU).  Runs the forward (or backward)
recurrence with DS2_RNN_PROF=1 and prints the mean clock-cycle gap between the in-kernel stamps of CTA (0,0,0)."""
import os
import sys

os.environ["DS2_RNN_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepspeech_b200 import ops
from deepspeech_b200._lib import DS2_BF16, check, lib

T, B, H = int(sys.argv[1]) if len(sys.argv) > 1 else 366, 32, int(sys.argv[2]) if len(sys.argv) > 2 else 1760
dev = "cuda"
wx = torch.randn(T * B, 2 * H, device=dev) * 0.5
U = (torch.randn(2, H, H, device=dev) * (0.5 / H ** 0.5)).to(torch.bfloat16)
bias = torch.zeros(2, H, device=dev)
lens = torch.full((B,), T, dtype=torch.int32, device=dev)
hs = torch.empty(2, T, B, H, dtype=torch.bfloat16, device=dev)
wsb = lib.ds2_rnn_workspace_bytes(T, B, H, 2, DS2_BF16)
ws = torch.zeros(wsb, dtype=torch.uint8, device=dev)
p, st = ops._p, ops._stream
for it in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(lib.ds2_rnn_relu_fwd(p(wx), 1.0, p(U), p(bias), p(lens), p(hs), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "fwd")
    e1.record()
    torch.cuda.synchronize()
print("fwd kernel %.3f ms  (%.2f us/step)" % (e0.elapsed_time(e1), e0.elapsed_time(e1) * 1e3 / T))
stamps = ws[1024:1024 + min(T, 1024) * 64].view(torch.int64).cpu().numpy().reshape(-1, 8)
s = stamps[2:min(T, 1024) - 1]
nxt = stamps[3:min(T, 1024)]
names = ["poll->tma_landed", "tma_landed->mma_issued", "mma_issued->mma_done(epi)", "mma_done->scatter_done", "scatter->cluster_barrier",
         "cluster_barrier->stores_issued", "stores->published", "published->next poll done"]
gaps = [s[:, 1] - s[:, 0], s[:, 2] - s[:, 1], s[:, 3] - s[:, 2], s[:, 4] - s[:, 3], s[:, 5] - s[:, 4], s[:, 6] - s[:, 5], s[:, 7] - s[:, 6],
        nxt[:, 0] - s[:, 7]]
tot = 0
for n, g in zip(names, gaps):
    print("%-34s mean %8.1f cyc   median %8.1f   p90 %8.1f" % (n, g.mean(), np.median(g), np.percentile(g, 90)))
    tot += g.mean()
print("sum %.1f cycles per step" % tot)
