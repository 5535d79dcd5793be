"""Allocation-free CTC entry point (ds2_ctc_loss_fwd_bwd_noalloc) against the regular one.

Who needs it: code that calls this library from a place where it cannot get scratch memory - a framework
whose own arena already owns all of HBM (cudaMalloc would fail), or a step loop that must not pay
cudaMalloc/cudaFree's device stall.  The variant keeps its alpha/beta tables in a 64 MiB array built into the
library, so every call shares that array and sees whatever the previous call left in it.  The test therefore
makes two calls in a row with different problem sizes (large first, then small and ragged) and requires each to
match ds2_ctc_loss_fwd_bwd - run with a caller-provided workspace - to the tolerance the other CTC tests use
(loss 1e-4 relative, gradient 1e-3 of its scale), and checks that a problem larger than the built-in array is
refused with a clear message.  Collected last (file name) so a -x run has already exercised everything else."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(T, B, L, ragged, seed):
    rng = np.random.default_rng(seed)
    x = torch.tensor(rng.standard_normal((T, B, 29)), dtype=torch.float32, device="cuda")
    labs = []
    for _ in range(B):
        n = int(rng.integers(max(1, L // 2), L + 1)) if ragged else L
        r = rng.integers(0, 28, size=n)
        labs.append([int(v) for i, v in enumerate(r) if i == 0 or v != r[i - 1]])
    Lmax = max(len(l) for l in labs)
    lab = np.zeros((B, Lmax), np.int32)
    for b, l in enumerate(labs):
        lab[b, :len(l)] = l
    ll = np.asarray([len(l) for l in labs], np.int32)
    sl = (rng.integers(max(T // 2, Lmax + 1), T + 1, size=B) if ragged else np.full(B, T)).astype(np.int32)
    dev = lambda a: torch.from_numpy(a).cuda()
    return x, dev(lab), dev(ll), dev(sl), Lmax


def _both(x, lab, ll, sl, Lmax):
    from deepspeech_b200 import _lib
    lib = _lib.lib
    T, B, K = x.shape
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    n = lib.ds2_ctc_workspace_bytes(T, B, K, Lmax)
    ws = torch.empty(n, dtype=torch.uint8, device="cuda")
    loss_a, grad_a = torch.empty(B, device="cuda"), torch.empty_like(x)
    _lib.check(lib.ds2_ctc_loss_fwd_bwd(P(x), T, B, K, P(lab), Lmax, P(ll), P(sl), None, P(loss_a), P(grad_a), P(ws), n, st), "regular")
    loss_b, grad_b = torch.full((B,), float("nan"), device="cuda"), torch.full_like(x, float("nan"))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]            # cudaMemGetInfo: what the DRIVER has left, not torch's cache
    _lib.check(lib.ds2_ctc_loss_fwd_bwd_noalloc(P(x), T, B, K, P(lab), Lmax, P(ll), P(sl), None, P(loss_b), P(grad_b), st), "noalloc")
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the allocation-free entry point changed the free device memory"
    return loss_a.cpu().numpy(), grad_a.cpu().numpy(), loss_b.cpu().numpy(), grad_b.cpu().numpy()


def test_noalloc_matches_regular_entry_point_on_two_calls_in_a_row():
    from deepspeech_b200 import _lib
    from oracle import ds2_oracle as O
    # the set-up call comes first, as a caller with its own arena would make it (before that arena claims HBM)
    assert _lib.lib.ds2_ctc_noalloc_max_bytes() == 64 << 20
    for (T, B, L, ragged, seed) in [(366, 32, 209, False, 1), (90, 5, 17, True, 2)]:
        x, lab, ll, sl, Lmax = _case(T, B, L, ragged, seed)
        la, ga, lb, gb = _both(x, lab, ll, sl, Lmax)
        assert np.isfinite(lb).all() and np.isfinite(gb).all()
        assert np.allclose(lb, la, rtol=1e-4, atol=1e-4)
        assert np.abs(gb - ga).max() <= 1e-3 * np.abs(ga).max()
        # and against the fp64 oracle on the first rows (not only CUDA against CUDA)
        nb = min(B, 3)
        labs = [[int(v) for v in lab[b, :int(ll[b])].cpu().numpy()] for b in range(nb)]
        lo, go = O.ctc_loss_and_grad_np(x[:, :nb].double().cpu().numpy(), labs, sl[:nb].cpu().numpy())
        assert np.abs(lb[:nb] - lo).max() <= 1e-4 * np.abs(lo).max()
        assert np.abs(gb[:, :nb] - go).max() <= 1e-3 * np.abs(go).max()


def test_noalloc_refuses_problems_beyond_its_built_in_scratch():
    from deepspeech_b200 import _lib
    x, lab, ll, sl, Lmax = _case(1600, 64, 400, False, 3)
    T, B, K = x.shape
    assert _lib.lib.ds2_ctc_workspace_bytes(T, B, K, Lmax) > _lib.lib.ds2_ctc_noalloc_max_bytes()
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    loss, grad = torch.empty(B, device="cuda"), torch.empty_like(x)
    rc = _lib.lib.ds2_ctc_loss_fwd_bwd_noalloc(P(x), T, B, K, P(lab), Lmax, P(ll), P(sl), None, P(loss), P(grad), None)
    assert rc == 3 and "built-in array holds" in _lib.last_error()
