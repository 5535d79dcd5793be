#!/usr/bin/env python
"""Timing of the implicit-GEMM conv2 kernels at the bench shape (needs a GPU).  usage: tools/conv_bench.py [B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deepspeech_b200 import ops
from deepspeech_b200._lib import check, lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
Ti, Fi, KT, KF, st, C = 741, 79, 10, 5, 2, 32
To, Fo = (Ti - KT) // st + 1, Fi - KF + 1
dev = "cuda"
a = ops._slack_bf16((B, Ti, Fi, C), dev)
a.copy_(torch.randn(B, Ti, Fi, C, device=dev))
w = torch.randn(KT, KF, C, C, device=dev) * 0.05
bias = torch.zeros(C, device=dev)
K = KT * KF * C
wt = torch.empty((C, K), dtype=torch.bfloat16, device=dev)
wd = torch.empty((C, K), dtype=torch.bfloat16, device=dev)
p, s = ops._p, ops._stream
check(lib.ds2_conv_window_weights(p(w), p(wt), p(wd), KT, KF, s()), "w")
y = torch.empty((B, To, Fo, C), device=dev)
dy = torch.randn(B, To, Fo, C, device=dev)
dyp = ops.conv_window_padded_dy(dy, B, Ti, Fi, KT, KF, st)
dw = torch.empty((KT, KF, C, C), device=dev)
da = torch.empty((B, Ti, Fi, C), device=dev)


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


flops = 2.0 * B * To * Fo * K * C
for name, fn in (("fwd", lambda: check(lib.ds2_conv_window_fwd(p(a), p(wt), p(bias), p(y), None, B, Ti, Fi, KT, KF, st, s()), "f")),
                 ("dgrad", lambda: check(lib.ds2_conv_window_dgrad(p(dyp), p(wd), p(da), B, Ti, Fi, KT, KF, st, s()), "d")),
                 ("wgrad", lambda: check(lib.ds2_conv_window_wgrad(p(a), p(dyp), p(dw), B, Ti, Fi, KT, KF, st, s()), "w"))):
    ms = timeit(fn)
    print("conv2 %-5s B=%d: %.3f ms  (%.0f TFLOP/s of the %.0f GFLOP)" % (name, B, ms, flops / ms / 1e9, flops / 1e9))

# conv1 (one input channel, filter [20,5], strides (2,2)) through the fused kernels
Ti1, Fi1 = 1500, 161
To1, Fo1 = (Ti1 - 20) // 2 + 1, (Fi1 - 5) // 2 + 1
x1 = torch.randn(B, Ti1, Fi1, device=dev)
w1 = torch.randn(20, 5, 1, C, device=dev) * 0.1
y1 = torch.empty(B, To1, Fo1, C, device=dev)
dy1 = torch.randn(B, To1, Fo1, C, device=dev)
sums = torch.empty(64, dtype=torch.float64, device=dev)
dw1 = torch.empty(20, 5, 1, C, device=dev)
db1 = torch.empty(C, device=dev)
mb = B * To1 * Fo1 * C * 4 / 1e6
for name, fn in (("fwd", lambda: check(lib.ds2_conv1_fused_fwd(p(x1), p(w1), p(bias), p(y1), p(sums), B, Ti1, Fi1, 20, 5, 2, 2, s()), "f")),
                 ("wgrad", lambda: check(lib.ds2_conv1_fused_wgrad(p(x1), p(dy1), p(dw1), p(db1), B, Ti1, Fi1, 20, 5, 2, 2, s()), "w"))):
    ms = timeit(fn)
    print("conv1 %-5s B=%d: %.3f ms  (%.0f MB of y / dy -> %.2f TB/s)" % (name, B, ms, mb, mb / ms / 1e3))
