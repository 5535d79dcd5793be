"""Worker of tests/test_gpu_dp.py (one process per GPU under torch.distributed.run, NCCL).
Every rank: (1) trains one step on ITS shard through GradAllReducer (gradients all-reduced from inside backward,
scaled 1/world); (2) computes, alone and without a reducer, the gradients of EVERY shard on its own GPU and averages them -
what the reference's sketched tower averaging (src/deepSpeech_train.py:193-228) computes, batch-norm statistics per tower;
(3) compares.  Prints 'DP_PARITY_OK <max relative deviation>' on success."""
import os
import sys
from types import SimpleNamespace

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)

from deepspeech_b200 import deepSpeech_NCHW as ds
from deepspeech_b200.dp import GradAllReducer, shard_batch
from oracle import ds2_oracle as O

use_bf16 = len(sys.argv) > 1 and sys.argv[1] == "bf16"
Bs, H, L = 16, 64, 2                                  # per-rank shard
params = SimpleNamespace(use_fp16=use_bf16, num_filters=32 if use_bf16 else 4, batch_size=Bs, num_hidden=H, num_rnn_layers=L,
                         engine="mkl", rnn_type="bi-dir", wiring="stacked", seed=7 + rank)      # different seeds: the reducer
m = ds.model(params)                                                                           # must make the replicas equal
red = GradAllReducer(m.parameters())
rng = np.random.default_rng(5)
feats, idx, vals, shape, seq_lens = O.dummy_batch(rng, Bs * world, 1)                           # T = 200 frames, global batch
labels_all = O.sparse_to_label_seqs(idx, vals, shape)


def sparse_of(rows):
    ii, vv = [], []
    for r, b in enumerate(rows):
        for c, v in enumerate(labels_all[b]):
            ii.append((r, c)); vv.append(v)
    Lm = max(len(labels_all[b]) for b in rows)
    return np.asarray(ii, np.int64).reshape(-1, 2), np.asarray(vv, np.int32), np.asarray([len(rows), Lm], np.int64)


def grads_of(rows, reducer):
    m.zero_grad(set_to_none=True)
    f = torch.from_numpy(feats[rows]).to(dev)
    logits = ds.inference(None, f, seq_lens[rows], params)
    loss = ds.loss(logits, sparse_of(rows), seq_lens[rows])
    loss.backward()
    if reducer is not None:
        reducer.finish()
    torch.cuda.synchronize()
    return [p.grad.detach().clone() for p in m.parameters()], float(loss.item())


sl = shard_batch(Bs * world, rank, world)
mine = list(range(sl.start, sl.stop))
g_dp, loss_dp = grads_of(mine, red)
red.remove()
ref = None
for r in range(world):
    s = shard_batch(Bs * world, r, world)
    g, _ = grads_of(list(range(s.start, s.stop)), None)
    ref = g if ref is None else [a + b for a, b in zip(ref, g)]
ref = [a / world for a in ref]
# bf16 path: split-K atomics reorder fp32 sums from run to run.  fp32 path: fp32 round-off - its weight-gradient and batch-norm
# reductions are atomic sums too (6e-6 is typical, 1e-5 has been seen once in a full-suite run)
tol = 2e-2 if use_bf16 else 5e-5
worst = 0.0
gmax = max(float(b.abs().max()) for b in ref)
for (name, _), a, b in zip(m.named_parameters(), g_dp, ref):
    # relative to the tensor's own scale, with a floor: the gradient of a bias that batch norm cancels is pure round-off
    e = float((a - b).abs().max() / max(float(b.abs().max()), 1e-3 * gmax))
    worst = max(worst, e)
    assert e < tol, (name, e)
# every rank holds the same reduced gradients
for a in g_dp:
    lo, hi = a.clone(), a.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert torch.equal(lo, hi)
if rank == 0:
    print("DP_PARITY_OK %.2e (loss %.5f, %d all-reduce calls)" % (worst, loss_dp, red.launched), flush=True)
dist.destroy_process_group()
