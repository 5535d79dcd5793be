"""CPU-only tests: the C-ABI library loads and exports every symbol include/ds2_b200.h declares,
argument validation works without a GPU, and the host-side logic (label transform, lengths, dummy
generator, initialisers, DP sharding + gradient all-reduce over gloo, world_size 2) is right."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import ds2_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    from deepspeech_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "ds2_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(ds2_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 25
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), "library does not export %s" % name
    assert sorted(_lib.EXPORTS) == declared, set(_lib.EXPORTS) ^ set(declared)
    assert _lib.lib.ds2_version() == 200


def test_c_abi_argument_validation_without_gpu():
    from deepspeech_b200 import _lib
    lib = _lib.lib
    assert lib.ds2_ctc_workspace_bytes(0, 1, 29, 1) == 0
    n = lib.ds2_ctc_workspace_bytes(366, 32, 29, 209)
    assert n >= 2 * 4 * 32 * 366 * 419
    rc = lib.ds2_ctc_loss_fwd_bwd(None, 10, 2, 29, None, 3, None, None, None, None, None, None, 0, None)
    assert rc == 1 and "null" in _lib.last_error()
    with pytest.raises(ValueError):
        _lib.check(rc, "ds2_ctc_loss_fwd_bwd")
    rc = lib.ds2_gemm_bf16(None, 8, 0, None, 8, 0, None, 0, 8, 4, 4, 8, 1.0, 0.0, None, None)
    assert rc == 1
    rc = lib.ds2_rnn_relu_fwd(None, 1.0, None, None, None, None, 0, 4, 2, 8, 2, 2, 20.0, None, 0, None)
    assert rc == 1
    with pytest.raises(RuntimeError):
        _lib.check(4, "x")
    # shape queries and host-only state of the newer entry points (no kernel is launched)
    assert lib.ds2_rnn_relu_persistent_supported(32, 1760, 2) == 1 and lib.ds2_rnn_relu_persistent_supported(33, 1760, 2) == 0
    assert lib.ds2_gru_persistent_supported(32, 800, 2) == 1 and lib.ds2_gru_persistent_supported(64, 1760, 2) == 1
    assert lib.ds2_gru_persistent_supported(48, 800, 2) == 0            # more than one chunk only in whole chunks of 32
    assert lib.ds2_conv_window_supported(32, 32, 79, 10, 5, 1) == 1 and lib.ds2_conv_window_supported(1, 32, 161, 20, 5, 2) == 0
    assert lib.ds2_set_sm_limit(36) == 0 and lib.ds2_set_sm_limit(0) == 36 and lib.ds2_set_sm_limit(0) == 0
    assert lib.ds2_adam_ema_chunk() > 0
    assert lib.ds2_adam_ema_step(None, 1, 1, 1e-3, 0.9, 0.999, 1e-8, 1, 0.99, 1.0, None) == 1
    assert lib.ds2_ctc_greedy_decode(None, 4, 1, 29, None, None, None, None) == 1
    assert lib.ds2_conv_window_fwd(None, None, None, None, None, 1, 20, 20, 3, 3, 1, None) == 1


def test_no_cpu_fallback():
    from deepspeech_b200 import ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.relux(torch.zeros(4))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.ctc_loss_dense(torch.zeros(4, 1, 29), np.zeros((1, 1), np.int32), np.zeros(1, np.int32), np.zeros(1, np.int32), [4])
    import deepspeech_b200
    src = "".join(open(os.path.join(ROOT, "deepspeech_b200", f)).read() for f in os.listdir(os.path.join(ROOT, "deepspeech_b200"))
                  if f.endswith(".py"))
    assert "oracle" not in src.replace("load_oracle_params", "").replace("oracle_grads", "").replace("oracle's", "") \
        .replace("from the oracle", "").replace("the oracle", ""), "product code must not import the oracle"


def test_label_transform_matches_oracle():
    from deepspeech_b200.ops import ctc_prepare_labels
    rng = np.random.default_rng(0)
    for _ in range(20):
        B, L = int(rng.integers(1, 6)), int(rng.integers(1, 12))
        dense = rng.integers(0, 28, size=(B, L))
        dense[:, -1] = 28
        idx = np.argwhere(dense != 0)
        vals = dense[dense != 0]
        for collapse in (True, False):
            seqs = O.sparse_to_label_seqs(idx, vals, dense.shape, 29, collapse)
            lab, ll, reps = ctc_prepare_labels(idx, vals, dense.shape, 29, collapse)
            assert [list(lab[b, :ll[b]]) for b in range(B)] == seqs
            assert all(reps[b] == sum(1 for i in range(1, len(s)) if s[i] == s[i - 1]) for b, s in enumerate(seqs))
    with pytest.raises(ValueError):
        ctc_prepare_labels([[0, 0], [0, 1]], [28, 3], [1, 2], 29, True)
    lab, ll, _ = ctc_prepare_labels(np.zeros((0, 2)), np.zeros(0), [2, 3], 29, True)     # empty labels
    assert ll.tolist() == [0, 0]


def test_lengths_and_dummy_generator():
    from deepspeech_b200.deepSpeech_dummy import DummyData, label_lengths, utt_lengths
    from deepspeech_b200.model import get_rnn_seqlen
    for T in list(range(39, 200)) + [1500, 6034]:
        assert get_rnn_seqlen([T])[0] == O.get_rnn_seqlen([T])[0] == O.conv_out_lens(T)[1]
    d = DummyData(4, seed=3)
    seen = []
    for _ in range(3):
        feats, idx, vals, shape, lens = d.inputs()
        seen.append(feats.shape[1])
        assert feats.dtype == np.float32 and feats.shape == (4, lens[0], 161) and (lens == lens[0]).all()
        assert (vals != 0).all() and shape[0] == 4
        seqs = O.sparse_to_label_seqs(idx, vals, shape)
        assert all(s == seqs[0] for s in seqs)
    assert seen == [100, 100, 100]
    f, idx, vals, shape, lens = d.inputs(bucket=14)
    assert f.shape == (4, utt_lengths[14], 161) and shape[1] == label_lengths[14]


def test_initialisers():
    from deepspeech_b200.helper_routines import glorot_uniform, variance_scaling
    rng = np.random.default_rng(0)
    w = variance_scaling(rng, (10, 5, 32, 32))
    assert abs(w.std() / np.sqrt(1.3 * 2 / 1600) - 0.88) < 0.02 and np.abs(w).max() <= 2 * np.sqrt(1.3 * 2 / 1600) + 1e-6
    u = glorot_uniform(rng, (64, 96))
    assert np.abs(u).max() <= np.sqrt(6 / 160) and u.shape == (64, 96)


def _dp_worker(rank, world, port, q):
    import torch.distributed as dist
    from deepspeech_b200.dp import GradAllReducer, shard_batch
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 8), torch.nn.ReLU(), torch.nn.Linear(8, 3)).double()
    X = torch.randn(8, 6, dtype=torch.float64)
    Y = torch.randn(8, 3, dtype=torch.float64)
    red = GradAllReducer(model.parameters(), bucket_bytes=200)
    sl = shard_batch(8, rank, world)
    loss = ((model(X[sl]) - Y[sl]) ** 2).mean()
    loss.backward()
    red.finish()
    g_dp = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    model.zero_grad()
    red.remove()
    ((model(X) - Y) ** 2).mean().backward()                # single-process global-batch gradient
    g_ref = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    q.put((rank, float((g_dp - g_ref).abs().max()), red.launched))
    dist.destroy_process_group()


def _free_port():
    """A TCP port nobody listens on right now (asked from the kernel), for the gloo rendezvous of the two workers."""
    import socket
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as sk:
        sk.bind(("127.0.0.1", 0))
        return sk.getsockname()[1]


class _EarlyLinear(torch.autograd.Function):
    """Stands in for the recurrent stack's autograd node: offers its weight gradient to ops.STACK_GRAD_HOOK from inside
    backward, as deepspeech_b200.ops._RecStackFn does layer by layer."""

    @staticmethod
    def forward(ctx, x, W):
        ctx.save_for_backward(x, W)
        ctx.param_objs = (W,)
        return x @ W.t()

    @staticmethod
    def backward(ctx, go):
        from deepspeech_b200 import ops
        x, W = ctx.saved_tensors
        gW = go.t() @ x
        if ops.STACK_GRAD_HOOK is not None:
            ops.STACK_GRAD_HOOK(0, [gW], ctx.param_objs)
        return go @ W, gW


def _dp_early_worker(rank, world, port, q):
    import torch.distributed as dist
    from deepspeech_b200 import ops
    from deepspeech_b200.dp import GradAllReducer, shard_batch
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    torch.manual_seed(0)
    W1 = torch.nn.Parameter(torch.randn(8, 6, dtype=torch.float64))
    lin2 = torch.nn.Linear(8, 3).double()
    X = torch.randn(8, 6, dtype=torch.float64)
    Y = torch.randn(8, 3, dtype=torch.float64)
    params = [W1] + list(lin2.parameters())
    fwd = lambda x: lin2(torch.relu(_EarlyLinear.apply(x, W1)))
    red = GradAllReducer(params, bucket_bytes=200)
    assert ops.STACK_GRAD_HOOK is not None
    sl = shard_batch(8, rank, world)
    errs = []
    for it in range(2):                                      # twice: the reducer must be reusable step after step
        for p in params:
            p.grad = None
        ((fwd(X[sl]) - Y[sl]) ** 2).mean().backward()
        red.finish()
        g_dp = torch.cat([p.grad.reshape(-1) for p in params])
        for p in params:
            p.grad = None
        hook, ops.STACK_GRAD_HOOK = ops.STACK_GRAD_HOOK, None
        ((fwd(X) - Y) ** 2).mean().backward()                # single-process global-batch gradient
        ops.STACK_GRAD_HOOK = hook
        g_ref = torch.cat([p.grad.reshape(-1) for p in params])
        errs.append(float((g_dp - g_ref).abs().max()))
    # gradients that were not reset: refused instead of silently double counting
    refused = False
    try:
        ((fwd(X[sl]) - Y[sl]) ** 2).mean().backward()
    except RuntimeError:
        refused = True
    red.remove()
    q.put((rank, max(errs), refused and ops.STACK_GRAD_HOOK is None))
    dist.destroy_process_group()


def test_dp_early_allreduce_from_stack_hook_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_early_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(err < 1e-12 and ok for _, err, ok in res), res


def test_dp_allreduce_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(err < 1e-12 for _, err, _ in res), res
    assert all(n >= 2 for _, _, n in res)                  # buckets were actually launched during backward


def test_shard_batch():
    from deepspeech_b200.dp import shard_batch
    assert [shard_batch(512, r, 8) for r in (0, 7)] == [slice(0, 64), slice(448, 512)]
    with pytest.raises(ValueError):
        shard_batch(10, 0, 4)
