#!/usr/bin/env python
"""Soak test (needs a GPU): many launches of the persistent recurrence kernels, checking that every run is bit-identical
to the first (the fence-free exchange must never let a stale row through), and many training steps checking the loss.
usage: tools/soak.py [minutes]"""
import os
import sys
import time
from types import SimpleNamespace

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepspeech_b200 import deepSpeech_NCHW as ds
from deepspeech_b200 import ops
from deepspeech_b200._lib import DS2_BF16, check, lib
from deepspeech_b200.deepSpeech_dummy import DummyData

minutes = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
dev = "cuda"
p, st = ops._p, ops._stream
g = torch.Generator(device=dev).manual_seed(3)
t_end = time.time() + minutes * 60.0 * 0.5
n = 0
for (T, B, H) in [(366, 32, 1760), (1500, 32, 1760), (366, 64, 800), (200, 16, 328)]:
    wx = torch.randn((T, B, 2 * H), device=dev, generator=g) * 0.7
    U = (torch.randn((2, H, H), device=dev, generator=g) * (0.6 / H ** 0.5)).to(torch.bfloat16)
    bias = torch.randn((2, H), device=dev, generator=g) * 0.1
    lens = torch.randint(T // 2, T + 1, (B,), device=dev, generator=g, dtype=torch.int32)
    dout = torch.randn((T, B, H), device=dev, generator=g)
    wsb = lib.ds2_rnn_workspace_bytes(T, B, H, 2, DS2_BF16)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    ref = None
    t_shape = time.time() + minutes * 60.0 * 0.5 / 4
    while time.time() < t_shape:
        hs = torch.empty((2, T, B, H), dtype=torch.bfloat16, device=dev)
        dz = torch.empty((T, B, 2 * H), dtype=torch.bfloat16, device=dev)
        check(lib.ds2_rnn_relu_fwd(p(wx), 1.0, p(U), p(bias), p(lens), p(hs), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "fwd")
        check(lib.ds2_rnn_relu_bwd(p(dout), p(U), p(lens), p(hs), p(dz), None, DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "bwd")
        torch.cuda.synchronize()
        if ref is None:
            ref = (hs, dz)
        else:
            assert torch.equal(hs, ref[0]) and torch.equal(dz, ref[1]), "run %d of T=%d B=%d H=%d differs from the first" % (n, T, B, H)
        n += 1
    print("recurrence T=%d B=%d H=%d: all runs bit-identical (%d launches so far)" % (T, B, H, 2 * n), flush=True)

# the GRU recurrence (per-warp publication since late round 2): same check
for (T, B, H) in [(366, 32, 800), (366, 64, 800), (200, 32, 1760)]:
    if not lib.ds2_gru_persistent_supported(B, H, 2):
        continue
    wx = torch.randn((T, B, 6 * H), device=dev, generator=g) * 0.5
    U = (torch.randn((2, 3 * H, H), device=dev, generator=g) * (0.5 / H ** 0.5)).to(torch.bfloat16)
    bw = torch.randn((2, 3 * H), device=dev, generator=g) * 0.1
    bu = torch.randn((2, 3 * H), device=dev, generator=g) * 0.1
    lens = torch.randint(T // 2, T + 1, (B,), device=dev, generator=g, dtype=torch.int32)
    dout = torch.randn((T, B, H), device=dev, generator=g)
    wsb = lib.ds2_gru_workspace_bytes(T, B, H, 2)
    ws = torch.zeros(wsb, dtype=torch.uint8, device=dev)
    ref, k = None, 0
    t_shape = time.time() + minutes * 60.0 * 0.25 / 3
    while time.time() < t_shape:
        hs = torch.empty((2, T, B, H), dtype=torch.bfloat16, device=dev)
        gates = torch.empty((2, T, B, 4 * H), dtype=torch.bfloat16, device=dev)
        dgi = torch.empty((T, B, 6 * H), dtype=torch.bfloat16, device=dev)
        dgh = torch.empty((T, B, 6 * H), dtype=torch.bfloat16, device=dev)
        check(lib.ds2_gru_persistent_fwd(p(wx), 1.0, p(U), p(bw), p(bu), p(lens), p(hs), p(gates), T, B, H, 2, 2, p(ws), wsb, st()), "gru fwd")
        check(lib.ds2_gru_persistent_bwd(p(dout), p(U), p(lens), p(hs), p(gates), p(dgi), p(dgh), None, None, T, B, H, 2, 2, p(ws), wsb, st()), "gru bwd")
        torch.cuda.synchronize()
        if ref is None:
            ref = (hs, gates, dgi, dgh)
        else:
            for a_, b_ in zip((hs, gates, dgi, dgh), ref):
                assert torch.equal(a_, b_), "GRU run %d of T=%d B=%d H=%d differs from the first" % (k, T, B, H)
        k += 1
    print("GRU recurrence T=%d B=%d H=%d: %d runs bit-identical" % (T, B, H, k), flush=True)

params = SimpleNamespace(use_fp16=True, num_filters=32, batch_size=32, num_hidden=1760, num_rnn_layers=7, engine="mkl",
                         rnn_type="bi-dir", wiring="stacked", seed=1234)
m = ds.model(params)
feats, idx, vals, shape, seq_lens = DummyData(32, seed=2).inputs(bucket=14)
x = torch.from_numpy(feats).cuda()
losses = []
t_end = time.time() + minutes * 60.0 * 0.5
while time.time() < t_end:
    m.zero_grad(set_to_none=True)
    loss = ds.loss(ds.inference(None, x, seq_lens, params), (idx, vals, shape), seq_lens)
    loss.backward()
    losses.append(loss.item())
    gn = float(sum(q.grad.double().pow(2).sum() for q in m.parameters()).sqrt())
    assert np.isfinite(losses[-1]) and np.isfinite(gn)
    if len(losses) > 1:
        assert abs(losses[-1] - losses[0]) <= 1e-5 * abs(losses[0]), (losses[0], losses[-1])
print("training step: %d steps, loss %.4f every time (forward is deterministic)" % (len(losses), losses[0]))
