#!/usr/bin/env python
"""Per-stage latency of the persistent recurrent kernel (needs a GPU).  Runs the forward recurrence with
DS2_RNN_PROF=1 and prints the mean clock-cycle gap between the in-kernel stamps of CTA (0,0,0), instance 0.
usage: tools/rnn_prof.py [T] [H] [B]"""
import os
import sys

os.environ["DS2_RNN_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepspeech_b200 import ops
from deepspeech_b200._lib import DS2_BF16, check, lib

T = int(sys.argv[1]) if len(sys.argv) > 1 else 366
H = int(sys.argv[2]) if len(sys.argv) > 2 else 1760
B = int(sys.argv[3]) if len(sys.argv) > 3 else 32
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
print("T=%d H=%d B=%d: fwd kernel %.3f ms  (%.2f us/step)" % (T, H, B, e0.elapsed_time(e1), e0.elapsed_time(e1) * 1e3 / T))
n = min(T, 1024)
S = ws[2048:2048 + n * 128].view(torch.int64).cpu().numpy().reshape(-1, 16)
s, nxt = S[10:n - 1], S[11:n]
if os.environ.get("DS2_RNN_V1"):
    seq = [("acquire counter + bar.sync", 8, 15), ("load h_{t-1} slice (7 x 16 B / thread)", 15, 9), ("swizzled st.shared", 9, 13), ("fence.proxy.async.shared::cta", 13, 14), ("mbarrier.arrive", 14, 10),
           ("loaders arrived -> control sees b_ready", 10, 0), ("MMAs + commit issued", 0, 1), ("commit -> epilogue sees accumulators", 1, 3),
           ("tcgen05.ld", 3, 11), ("st.async partial slabs issued", 11, 12), ("all 4 slabs arrived", 12, 5),
           ("reduce + epilogue + h_t store issued", 5, 6), ("bar.sync + red.release issued", 6, 7)]
    last, first = 7, 8
else:
    # round-2 protocol: control thread stamps 0, 1, 4; first epilogue thread 13, 8, 9..12
    seq = [("control: wait for the step counter", 0, 1), ("bulk copy landed (seen by the epilogue warp)", 1, 13),
           ("image checked + arrive", 13, 8), ("accumulators + verdict seen", 8, 9),
           ("tcgen05.ld + st.async partial slabs issued", 9, 10), ("all 4 slabs arrived", 10, 11),
           ("reduce + epilogue + stores + counter red", 11, 12)]
    last, first = 12, 0
tot = 0.0
for name, a, b in seq:
    g = (s[:, b] - s[:, a]).astype(np.float64)
    print("%-46s mean %8.1f cyc   median %8.1f   p90 %8.1f" % (name, g.mean(), np.median(g), np.percentile(g, 90)))
    tot += g.mean()
g = (nxt[:, first] - s[:, last]).astype(np.float64)
print("%-46s mean %8.1f cyc" % ("published -> next step starts (negative: the control thread is already waiting)", g.mean()))
print("sum %.0f cycles per step" % (tot + g.mean()))
if os.environ.get("DS2_RNN_V1"):
    r = s[:, 2].astype(np.float64)
    print("reloads of not-yet-visible rows by the stamping thread: mean %.2f per step, steps with any: %.0f %%" % (r.mean(), 100.0 * (r > 0).mean()))
else:
    g = (s[:, 4] - s[:, 1]).astype(np.float64)
    print("control: counter seen -> copy landed, MMAs issued, commit: mean %.1f cyc;  re-fetches: %d" % (g.mean(), int(S[:, 7].sum())))
    g = (s[:, 12] - s[:, 1]).astype(np.float64)
    print("counter seen -> own publication: mean %.1f cyc  (the rest of the step is the wait for the other CTAs)" % g.mean())
if not os.environ.get("DS2_RNN_V1"):
    # per-CTA totals (instance 0): time spent waiting for the other CTAs and in fetch + MMA issue, SM id, counter copy polled
    nct = 4 * ((H + 127) // 128) * 2
    C = ws[2048 + 8192 * 8:2048 + (8192 + nct * 4) * 8].view(torch.int64).cpu().numpy().reshape(-1, 4)
    steps = T - 1 - 8
    wait, fetch = C[:, 0] / steps, C[:, 1] / max(T - 1, 1)
    order = np.argsort(wait)
    print("per-CTA (instance 0): wait for the others / fetch+issue, clk per step; the CTAs that wait least set the pace")
    print("  wait: min %.0f  p10 %.0f  median %.0f  p90 %.0f  max %.0f   fetch+issue: min %.0f median %.0f max %.0f"
          % (wait.min(), np.percentile(wait, 10), np.median(wait), np.percentile(wait, 90), wait.max(), fetch.min(), np.median(fetch), fetch.max()))
    for i in list(order[:8]) + list(order[-4:]):
        d, rest = divmod(int(i), nct // 2)
        print("   dir %d tile %2d rank %d  sm %3d  copy %d: wait %5.0f  fetch+issue %5.0f" % (d, rest // 4, rest % 4, C[i, 2], C[i, 3], wait[i], fetch[i]))
    print("  counter copies chosen: %s" % np.bincount(C[:, 3].astype(int), minlength=4))
