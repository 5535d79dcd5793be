#!/usr/bin/env python
"""Event-timed persistent recurrence kernels, ReLU and GRU, forward and backward (needs a GPU).
usage: tools/rec_bench.py [H] [B] [T]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deepspeech_b200 import ops
from deepspeech_b200._lib import DS2_BF16, check, lib

H = int(sys.argv[1]) if len(sys.argv) > 1 else 800
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
T = int(sys.argv[3]) if len(sys.argv) > 3 else 366
dev = "cuda"
p, st = ops._p, ops._stream
lens = torch.full((B,), T, dtype=torch.int32, device=dev)
dout = torch.randn(T, B, H, device=dev)


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


if lib.ds2_rnn_relu_persistent_supported(B, H, 2):
    wx = torch.randn(T, B, 2 * H, device=dev) * 0.5
    U = (torch.randn(2, H, H, device=dev) * (0.5 / H ** 0.5)).to(torch.bfloat16)
    bias = torch.zeros(2, H, device=dev)
    hs = torch.empty(2, T, B, H, dtype=torch.bfloat16, device=dev)
    dz = torch.empty(T, B, 2 * H, dtype=torch.bfloat16, device=dev)
    db = torch.empty(2, H, device=dev)
    wsb = lib.ds2_rnn_workspace_bytes(T, B, H, 2, DS2_BF16)
    ws = torch.zeros(wsb, dtype=torch.uint8, device=dev)
    f = timeit(lambda: check(lib.ds2_rnn_relu_fwd(p(wx), 1.0, p(U), p(bias), p(lens), p(hs), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "f"))
    b = timeit(lambda: check(lib.ds2_rnn_relu_bwd(p(dout), p(U), p(lens), p(hs), p(dz), p(db), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "b"))
    print("ReLU H=%d B=%d T=%d: fwd %.3f ms (%.2f us/step)  bwd %.3f ms (%.2f us/step)" % (H, B, T, f, f * 1e3 / T, b, b * 1e3 / T))
if lib.ds2_gru_persistent_supported(B, H, 2):
    wx = torch.randn(T, B, 6 * H, device=dev) * 0.5
    U = (torch.randn(2, 3 * H, H, device=dev) * (0.5 / H ** 0.5)).to(torch.bfloat16)
    bw = torch.zeros(2, 3 * H, device=dev)
    bu = torch.zeros(2, 3 * H, device=dev)
    hs = torch.empty(2, T, B, H, dtype=torch.bfloat16, device=dev)
    gates = torch.empty(2, T, B, 4 * H, dtype=torch.bfloat16, device=dev)
    dgi = torch.empty(T, B, 6 * H, dtype=torch.bfloat16, device=dev)
    dgh = torch.empty(T, B, 6 * H, dtype=torch.bfloat16, device=dev)
    dbw = torch.empty(2, 3 * H, device=dev)
    dbu = torch.empty(2, 3 * H, device=dev)
    wsb = lib.ds2_gru_workspace_bytes(T, B, H, 2)
    ws = torch.zeros(wsb, dtype=torch.uint8, device=dev)
    f = timeit(lambda: check(lib.ds2_gru_persistent_fwd(p(wx), 1.0, p(U), p(bw), p(bu), p(lens), p(hs), p(gates), T, B, H, 2, 2, p(ws), wsb, st()), "f"))
    b = timeit(lambda: check(lib.ds2_gru_persistent_bwd(p(dout), p(U), p(lens), p(hs), p(gates), p(dgi), p(dgh), p(dbw), p(dbu), T, B, H, 2, 2,
                                                        p(ws), wsb, st()), "b"))
    print("GRU  H=%d B=%d T=%d: fwd %.3f ms (%.2f us/step)  bwd %.3f ms (%.2f us/step)" % (H, B, T, f, f * 1e3 / T, b, b * 1e3 / T))
