#!/usr/bin/env python
"""Per-stage cycles of the persistent GRU kernels (needs a GPU): DS2_RNN_PROF stamps of CTA (0,0,0), instance 0.
usage: tools/gru_prof.py [T] [H] [B]"""
import os
import sys

os.environ["DS2_RNN_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepspeech_b200 import ops
from deepspeech_b200._lib import check, lib

T = int(sys.argv[1]) if len(sys.argv) > 1 else 366
H = int(sys.argv[2]) if len(sys.argv) > 2 else 800
B = int(sys.argv[3]) if len(sys.argv) > 3 else 32
dev = "cuda"
p, st = ops._p, ops._stream
lens = torch.full((B,), T, dtype=torch.int32, device=dev)
wx = torch.randn(T, B, 6 * H, device=dev) * 0.5
U = (torch.randn(2, 3 * H, H, device=dev) * (0.5 / H ** 0.5)).to(torch.bfloat16)
bw = torch.zeros(2, 3 * H, device=dev)
bu = torch.zeros(2, 3 * H, device=dev)
hs = torch.empty(2, T, B, H, dtype=torch.bfloat16, device=dev)
gates = torch.empty(2, T, B, 4 * H, dtype=torch.bfloat16, device=dev)
dgi = torch.empty(T, B, 6 * H, dtype=torch.bfloat16, device=dev)
dgh = torch.empty(T, B, 6 * H, dtype=torch.bfloat16, device=dev)
dout = torch.randn(T, B, H, device=dev)
wsb = lib.ds2_gru_workspace_bytes(T, B, H, 2)
ws = torch.zeros(wsb, dtype=torch.uint8, device=dev)
names = ["loop top + prefetch -> counter seen (bar.sync)", "loads, stale check, st.shared, fence, arrive", "wait for the MMAs",
         "tcgen05.ld + st.async of the slabs", "all slabs arrived", "reduce + gate math + stores issued", "bar.sync + counter increment"]


def report(tag):
    torch.cuda.synchronize()
    n = min(T, 1024)
    S = ws[2048:2048 + n * 64].view(torch.int64).cpu().numpy().reshape(-1, 8)[2:n - 1]
    tot = 0.0
    print(tag)
    for i, nm in enumerate(names):
        g = (S[:, i + 1] - S[:, i]).astype(np.float64)
        tot += g.mean()
        print("  %-52s mean %7.1f  median %7.1f  p90 %7.1f" % (nm, g.mean(), np.median(g), np.percentile(g, 90)))
    print("  sum %.0f cycles per step (stamp to stamp, without the loop back-edge)" % tot)


for _ in range(2):
    check(lib.ds2_gru_persistent_fwd(p(wx), 1.0, p(U), p(bw), p(bu), p(lens), p(hs), p(gates), T, B, H, 2, 2, p(ws), wsb, st()), "f")
report("GRU forward  T=%d H=%d B=%d" % (T, H, B))
for _ in range(2):
    check(lib.ds2_gru_persistent_bwd(p(dout), p(U), p(lens), p(hs), p(gates), p(dgi), p(dgh), None, None, T, B, H, 2, 2, p(ws), wsb, st()), "b")
report("GRU backward T=%d H=%d B=%d" % (T, H, B))
