"""GPU parity tests, kernel level: every C-ABI entry point against the CPU oracle (oracle/ds2_oracle.py)
or a float64 restatement of the same formula.  Tolerances: fp32 kernels 1e-4 rel. on losses, 1e-3 on
gradients (BASELINE.json north_star); bf16 tensor-core GEMM checked on bf16-rounded inputs (fp32
accumulate) to 1e-3 of the result's scale."""
import numpy as np
import pytest
import torch

from oracle import ds2_oracle as O

pytestmark = pytest.mark.gpu


def _ops():
    from deepspeech_b200 import ops
    return ops


def relerr(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


# ----------------------------------------------------------------------------------------- CTC
def _ctc_case(T, B, L, ragged, seed, K=29):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((T, B, K)).astype(np.float32)
    if ragged:
        lens = rng.integers(T // 2, T + 1, size=B)
        labels = [list(rng.integers(0, K - 1, size=rng.integers(max(L // 2, 0), L + 1))) for _ in range(B)]
    else:
        lens = np.full(B, T)
        labels = [list(rng.integers(0, K - 1, size=L)) for _ in range(B)]
    # TF collapses repeats in the dense label (preprocess_collapse_repeated=True in the reference call)
    labels = [[v for i, v in enumerate(l) if i == 0 or v != l[i - 1]] for l in labels]
    return x, labels, lens.astype(np.int32)


def _run_ctc(x, labels, lens):
    ops = _ops()
    B = x.shape[1]
    Lmax = max([len(l) for l in labels] + [1])
    lab = np.zeros((B, Lmax), np.int32)
    ll = np.zeros(B, np.int32)
    for b, l in enumerate(labels):
        lab[b, :len(l)] = l
        ll[b] = len(l)
    reps = np.asarray([sum(1 for i in range(1, len(l)) if l[i] == l[i - 1]) for l in labels], np.int32)
    xt = torch.tensor(x, device="cuda", requires_grad=True)
    loss = ops.ctc_loss_dense(xt, lab, ll, reps, lens)
    loss.sum().backward()
    torch.cuda.synchronize()
    return loss.detach().cpu().numpy(), xt.grad.cpu().numpy()


@pytest.mark.parametrize("T,B,L,ragged", [(1, 1, 1, False), (5, 2, 2, False), (12, 3, 0, False), (30, 4, 9, True),
                                           (64, 7, 20, True), (200, 16, 50, False), (366, 8, 209, False),
                                           (400, 8, 100, True)])
def test_ctc_vs_oracle(T, B, L, ragged):
    x, labels, lens = _ctc_case(T, B, L, ragged, seed=T * 31 + B)
    ref_loss, ref_grad = O.ctc_loss_and_grad_np(x, labels, lens)
    loss, grad = _run_ctc(x, labels, lens)
    assert np.allclose(loss, ref_loss, rtol=1e-4, atol=1e-4), (loss, ref_loss)
    assert relerr(grad, ref_grad) < 1e-3
    assert np.abs(grad.sum(2)).max() < 1e-3           # rows sum to zero inside (to the gradient bar), are zero outside
    for b in range(B):
        assert np.abs(grad[lens[b]:, b]).max(initial=0.0) == 0.0


def test_ctc_repeated_labels_without_collapse_and_closed_form():
    ops = _ops()
    # uniform logits, no repeats: loss = T ln K - ln C(T+L, 2L)
    import math
    T, L = 12, 5
    loss, grad = _run_ctc(np.zeros((T, 1, 29), np.float32), [list(range(1, L + 1))], np.asarray([T], np.int32))
    assert abs(loss[0] - (T * math.log(29) - math.log(math.comb(T + L, 2 * L)))) < 1e-4
    # a label with an adjacent repeat needs a blank between: handled, and needs T >= L + repeats
    x = np.random.default_rng(0).standard_normal((6, 1, 29)).astype(np.float32)
    ref_loss, ref_grad = O.ctc_loss_and_grad_np(x, [[3, 3, 4]], [6])
    loss, grad = _run_ctc(x, [[3, 3, 4]], np.asarray([6], np.int32))
    assert np.allclose(loss, ref_loss, rtol=1e-4) and relerr(grad, ref_grad) < 1e-3
    with pytest.raises(ValueError, match="Not enough time"):
        _run_ctc(x[:3], [[3, 3, 4]], np.asarray([3], np.int32))
    del ops


def test_ctc_sparse_front_door_matches_oracle_transform():
    from deepspeech_b200.ctc import ctc_loss
    rng = np.random.default_rng(4)
    T, B = 40, 3
    x = rng.standard_normal((T, B, 29)).astype(np.float32)
    dense = np.asarray([[5, 5, 0, 7, 28, 0], [9, 0, 9, 9, 1, 28], [0, 0, 0, 0, 0, 28]])
    idx = np.argwhere(dense != 0)
    vals = dense[dense != 0]
    seqs = O.sparse_to_label_seqs(idx, vals, dense.shape)
    assert seqs == [[5, 7], [9, 1], []]
    ref_loss, _ = O.ctc_loss_and_grad_np(x, seqs, [T, T - 5, T - 9])
    got = ctc_loss((idx, vals, dense.shape), torch.tensor(x, device="cuda"), np.asarray([T, T - 5, T - 9], np.int32),
                   preprocess_collapse_repeated=True)
    assert np.allclose(got.cpu().numpy(), ref_loss, rtol=1e-4)
    with pytest.raises(ValueError, match="non-null label"):
        ctc_loss((np.asarray([[0, 0], [0, 1]]), np.asarray([28, 3]), (1, 2)), torch.tensor(x[:, :1], device="cuda"),
                 np.asarray([T], np.int32), preprocess_collapse_repeated=True)


# ----------------------------------------------------------------------------------------- GEMM
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 384, 512), (100, 29, 72), (130, 200, 264), (1, 8, 8),
                                    (517, 1760, 2400)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (True, False), (False, True), (True, True)])
def test_gemm_tcgen05_vs_fp64(M, N, K, a_mn, b_mn):
    ops = _ops()
    rng = np.random.default_rng(M + N + K)
    Mp, Np = (M + 7) // 8 * 8, (N + 7) // 8 * 8          # MN-major operands need a row pitch multiple of 8
    A = torch.tensor(rng.standard_normal((K, Mp) if a_mn else (M, K)), dtype=torch.bfloat16, device="cuda")
    B = torch.tensor(rng.standard_normal((K, Np) if b_mn else (N, K)), dtype=torch.bfloat16, device="cuda")
    bias = torch.tensor(rng.standard_normal(N), dtype=torch.float32, device="cuda")
    C0 = torch.tensor(rng.standard_normal((M, N)), dtype=torch.float32, device="cuda")
    C = C0.clone()
    out = ops.gemm_nt(A, B, a_mn=a_mn, b_mn=b_mn, M=M, N=N, K=K, out=C, alpha=0.5, beta=2.0, bias=bias)
    torch.cuda.synchronize()
    Ad = A.double().cpu().numpy()
    Bd = B.double().cpu().numpy()
    Am = Ad[:, :M].T if a_mn else Ad
    Bm = Bd[:, :N].T if b_mn else Bd
    ref = 0.5 * Am @ Bm.T + bias.double().cpu().numpy() + 2.0 * C0.double().cpu().numpy()
    assert relerr(out.cpu().numpy(), ref) < 1e-3
    # bf16 output, no epilogue extras
    out2 = ops.gemm_nt(A, B, a_mn=a_mn, b_mn=b_mn, M=M, N=N, K=K, out_dtype=torch.bfloat16)
    assert relerr(out2.float().cpu().numpy(), Am @ Bm.T) < 1e-2


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (True, True), (True, False)])
def test_gemm_generic_vs_fp64(a_mn, b_mn):
    ops = _ops()
    rng = np.random.default_rng(9)
    M, N, K = 37, 29, 53
    A = torch.tensor(rng.standard_normal((K, M) if a_mn else (M, K)), dtype=torch.float32, device="cuda")
    B = torch.tensor(rng.standard_normal((K, N) if b_mn else (N, K)), dtype=torch.float32, device="cuda")
    bias = torch.tensor(rng.standard_normal(N), dtype=torch.float32, device="cuda")
    out = ops.gemm_nt(A, B, a_mn=a_mn, b_mn=b_mn, alpha=1.5, bias=bias)
    Am = A.double().cpu().numpy().T if a_mn else A.double().cpu().numpy()
    Bm = B.double().cpu().numpy().T if b_mn else B.double().cpu().numpy()
    assert relerr(out.cpu().numpy(), 1.5 * Am @ Bm.T + bias.double().cpu().numpy()) < 1e-5


# ----------------------------------------------------------------------------------------- conv / BN
def test_conv2d_fwd_bwd_vs_torch_fp64():
    ops = _ops()
    rng = np.random.default_rng(11)
    for (B, Ti, Fi, Ci, Co, KT, KF, st, sf) in [(2, 45, 161, 1, 4, 20, 5, 2, 2), (2, 13, 79, 4, 4, 10, 5, 2, 1)]:
        x = torch.tensor(rng.standard_normal((B, Ti, Fi, Ci)), dtype=torch.float32, device="cuda", requires_grad=True)
        w = torch.tensor(rng.standard_normal((KT, KF, Ci, Co)) * 0.1, dtype=torch.float32, device="cuda", requires_grad=True)
        b = torch.tensor(rng.standard_normal(Co), dtype=torch.float32, device="cuda", requires_grad=True)
        y = ops.conv2d_nhwc(x, w, b, st, sf)
        g = torch.tensor(rng.standard_normal(tuple(y.shape)), dtype=torch.float32, device="cuda")
        (y * g).sum().backward()
        xr = x.detach().double().cpu().permute(0, 3, 1, 2).requires_grad_(True)
        wr = w.detach().double().cpu().requires_grad_(True)
        br = b.detach().double().cpu().requires_grad_(True)
        yr = torch.nn.functional.conv2d(xr, wr.permute(3, 2, 0, 1), br, stride=(st, sf))
        (yr * g.double().cpu().permute(0, 3, 1, 2)).sum().backward()
        assert relerr(y.detach().cpu().permute(0, 3, 1, 2).numpy(), yr.detach().numpy()) < 1e-5
        assert relerr(w.grad.cpu().numpy(), wr.grad.numpy()) < 1e-4
        assert relerr(b.grad.cpu().numpy(), br.grad.numpy()) < 1e-4
        assert relerr(x.grad.cpu().permute(0, 3, 1, 2).numpy(), xr.grad.numpy()) < 1e-4


def test_conv_gemm_route_vs_fp64():
    """im2col (+ fused BN1/clip) -> tcgen05 GEMM route of the front-end against an fp64 restatement."""
    ops = _ops()
    rng = np.random.default_rng(21)
    B, T, nf = 2, 61, 8
    feats = torch.tensor(rng.standard_normal((B, T, 161, 1)), dtype=torch.float32, device="cuda")
    w1 = torch.tensor(rng.standard_normal((20, 5, 1, nf)) * 0.15, dtype=torch.float32, device="cuda", requires_grad=True)
    b1 = torch.tensor(rng.standard_normal(nf) * 0.1, dtype=torch.float32, device="cuda", requires_grad=True)
    g1 = torch.tensor(rng.uniform(0.6, 1.4, nf), dtype=torch.float32, device="cuda", requires_grad=True)
    be1 = torch.tensor(rng.standard_normal(nf) * 0.3, dtype=torch.float32, device="cuda", requires_grad=True)
    w2 = torch.tensor(rng.standard_normal((10, 5, nf, nf)) * 0.08, dtype=torch.float32, device="cuda", requires_grad=True)
    b2 = torch.tensor(rng.standard_normal(nf) * 0.1, dtype=torch.float32, device="cuda", requires_grad=True)
    y1 = ops.conv2d_gemm(feats, w1, b1, 2, 2)
    y2 = ops.conv2d_gemm(y1, w2, b2, 2, 1, g1, be1, 1e-3, 20.0)
    g = torch.tensor(rng.standard_normal(tuple(y2.shape)), dtype=torch.float32, device="cuda")
    (y2 * g).sum().backward()
    torch.cuda.synchronize()
    leaves = [t.detach().double().cpu().requires_grad_(True) for t in (w1, b1, g1, be1, w2, b2)]
    W1, B1, G1, BE1, W2, B2 = leaves
    xr = feats.double().cpu().permute(0, 3, 1, 2)
    r1 = torch.nn.functional.conv2d(xr, W1.permute(3, 2, 0, 1), B1, stride=(2, 2))
    a1 = O.relux(O.batch_norm_train(r1, G1, BE1, 1e-3, 1), 20.0)
    r2 = torch.nn.functional.conv2d(a1, W2.permute(3, 2, 0, 1), B2, stride=(2, 1))
    (r2 * g.double().cpu().permute(0, 3, 1, 2)).sum().backward()
    assert relerr(y1.detach().cpu().permute(0, 3, 1, 2).numpy(), r1.detach().numpy()) < 1e-2
    assert relerr(y2.detach().cpu().permute(0, 3, 1, 2).numpy(), r2.detach().numpy()) < 2e-2

    def cos(a, b):
        a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
        return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-300))
    for name, got, ref in (("w2", w2, W2), ("b2", b2, B2), ("w1", w1, W1), ("g1", g1, G1), ("be1", be1, BE1)):
        assert cos(got.grad.cpu().numpy(), ref.grad.numpy()) > 0.995, name
        assert abs(np.linalg.norm(got.grad.cpu().numpy()) / np.linalg.norm(ref.grad.numpy()) - 1) < 3e-2, name


@pytest.mark.parametrize("B,Ti,Fi,KT,KF,st", [(2, 31, 79, 10, 5, 2), (1, 12, 9, 3, 2, 1), (3, 22, 128, 4, 7, 3), (2, 741, 79, 10, 5, 2),
                                              (1, 25, 40, 5, 3, 2), (2, 18, 30, 3, 4, 1), (1, 10, 16, 2, 1, 1), (1, 9, 8, 8, 8, 4)])
def test_conv_window_kernels_vs_fp64(B, Ti, Fi, KT, KF, st):
    """Implicit-GEMM conv (forward, data gradient, weight gradient through the strided TMA window view) against an fp64
    convolution of the SAME bf16-rounded operands: only fp32 accumulation order differs (<= 1e-4 of the scale)."""
    ops = _ops()
    from deepspeech_b200._lib import check, lib
    C = 32
    assert lib.ds2_conv_window_supported(C, C, Fi, KT, KF, 1) == 1
    rng = np.random.default_rng(Ti * 3 + KF)
    To, Fo = (Ti - KT) // st + 1, Fi - KF + 1
    bf = lambda t: t.to(torch.bfloat16).float()
    x = bf(torch.tensor(rng.standard_normal((B, Ti, Fi, C)), dtype=torch.float32, device="cuda"))
    w = bf(torch.tensor(rng.standard_normal((KT, KF, C, C)) * 0.1, dtype=torch.float32, device="cuda"))
    bias = torch.tensor(rng.standard_normal(C), dtype=torch.float32, device="cuda")
    dy = bf(torch.tensor(rng.standard_normal((B, To, Fo, C)), dtype=torch.float32, device="cuda"))
    p, st_ = ops._p, ops._stream
    a = ops._slack_bf16((B, Ti, Fi, C), "cuda")
    a.copy_(x)
    K = KT * KF * C
    wt = torch.empty((C, K), dtype=torch.bfloat16, device="cuda")
    wd = torch.empty((C, K), dtype=torch.bfloat16, device="cuda")
    check(lib.ds2_conv_window_weights(p(w), p(wt), p(wd), KT, KF, st_()), "weights")
    y = torch.full((B, To, Fo, C), float("nan"), dtype=torch.float32, device="cuda")
    sums = torch.full((64,), float("nan"), dtype=torch.float64, device="cuda")
    check(lib.ds2_conv_window_fwd(p(a), p(wt), p(bias), p(y), p(sums), B, Ti, Fi, KT, KF, st, st_()), "fwd")
    dyp = ops.conv_window_padded_dy(dy, B, Ti, Fi, KT, KF, st)
    dw = torch.full((KT, KF, C, C), float("nan"), dtype=torch.float32, device="cuda")
    check(lib.ds2_conv_window_wgrad(p(a), p(dyp), p(dw), B, Ti, Fi, KT, KF, st, st_()), "wgrad")
    da = torch.full((B, Ti, Fi, C), float("nan"), dtype=torch.float32, device="cuda")
    check(lib.ds2_conv_window_dgrad(p(dyp), p(wd), p(da), B, Ti, Fi, KT, KF, st, st_()), "dgrad")
    torch.cuda.synchronize()
    xr = x.double().cpu().permute(0, 3, 1, 2).requires_grad_(True)
    wr = w.double().cpu().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xr, wr.permute(3, 2, 0, 1), bias.double().cpu(), stride=(st, 1))
    (yr * dy.double().cpu().permute(0, 3, 1, 2)).sum().backward()
    assert relerr(y.cpu().permute(0, 3, 1, 2).numpy(), yr.detach().numpy()) < 1e-4
    assert relerr(dw.cpu().numpy(), wr.grad.numpy()) < 1e-4
    assert relerr(da.cpu().permute(0, 3, 1, 2).numpy(), xr.grad.numpy()) < 1e-4
    # batch-norm sums from the forward epilogue
    flat = yr.detach().permute(0, 2, 3, 1).reshape(-1, C)
    assert np.allclose(sums[:C].cpu().numpy(), flat.sum(0).numpy(), rtol=1e-4, atol=1e-5 * float(flat.abs().sum(0).max()))
    assert np.allclose(sums[C:].cpu().numpy(), (flat * flat).sum(0).numpy(), rtol=1e-4)


@pytest.mark.parametrize("B,Ti,Fi,KT,KF,st,sf", [(2, 60, 161, 20, 5, 2, 2), (1, 25, 40, 5, 3, 2, 1), (3, 23, 17, 4, 8, 1, 2), (2, 300, 161, 20, 5, 2, 2),
                                                 (1, 9, 9, 9, 1, 3, 3), (2, 41, 12, 11, 2, 2, 1)])
def test_conv1_fused_vs_fp64(B, Ti, Fi, KT, KF, st, sf):
    """Fused one-channel convolution (patch rows built in shared memory, BN sums from the epilogue, weight + bias gradient)
    against an fp64 convolution of the SAME bf16-rounded operands: only fp32 accumulation order differs."""
    ops = _ops()
    from deepspeech_b200._lib import check, lib
    C = 32
    assert lib.ds2_conv1_fused_supported(1, C, Ti, Fi, KT, KF, st, sf) == 1
    rng = np.random.default_rng(Ti * 7 + KF)
    To, Fo = (Ti - KT) // st + 1, (Fi - KF) // sf + 1
    bf = lambda t: t.to(torch.bfloat16).float()
    x = bf(torch.tensor(rng.standard_normal((B, Ti, Fi)), dtype=torch.float32, device="cuda"))
    w = bf(torch.tensor(rng.standard_normal((KT, KF, 1, C)) * 0.2, dtype=torch.float32, device="cuda"))
    bias = torch.tensor(rng.standard_normal(C), dtype=torch.float32, device="cuda")
    dy = bf(torch.tensor(rng.standard_normal((B, To, Fo, C)), dtype=torch.float32, device="cuda"))
    p, st_ = ops._p, ops._stream
    y = torch.full((B, To, Fo, C), float("nan"), dtype=torch.float32, device="cuda")
    sums = torch.full((2 * C,), float("nan"), dtype=torch.float64, device="cuda")
    check(lib.ds2_conv1_fused_fwd(p(x), p(w), p(bias), p(y), p(sums), B, Ti, Fi, KT, KF, st, sf, st_()), "fwd")
    dw = torch.full((KT, KF, 1, C), float("nan"), dtype=torch.float32, device="cuda")
    db = torch.full((C,), float("nan"), dtype=torch.float32, device="cuda")
    check(lib.ds2_conv1_fused_wgrad(p(x), p(dy), p(dw), p(db), B, Ti, Fi, KT, KF, st, sf, st_()), "wgrad")
    mean = torch.empty(C, device="cuda")
    rstd = torch.empty(C, device="cuda")
    check(lib.ds2_bn_finalize(p(sums), float(B * To * Fo), C, 1e-3, p(mean), p(rstd), st_()), "finalize")
    torch.cuda.synchronize()
    xr = x.double().cpu().view(B, 1, Ti, Fi)
    wr = w.double().cpu().requires_grad_(True)
    br = bias.double().cpu().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xr, wr.permute(3, 2, 0, 1), br, stride=(st, sf))
    (yr * dy.double().cpu().permute(0, 3, 1, 2)).sum().backward()
    yn = yr.detach().permute(0, 2, 3, 1)
    assert relerr(y.cpu().numpy(), yn.numpy()) < 1e-4
    assert relerr(dw.cpu().numpy(), wr.grad.numpy()) < 1e-4
    assert relerr(db.cpu().numpy(), br.grad.numpy()) < 1e-4
    flat = yn.reshape(-1, C)
    assert np.allclose(mean.cpu().numpy(), flat.mean(0).numpy(), rtol=1e-4, atol=1e-4)
    assert np.allclose(rstd.cpu().numpy(), (1.0 / torch.sqrt(flat.var(0, unbiased=False) + 1e-3)).numpy(), rtol=1e-4)


def test_conv_window_route_matches_patch_matrix_route():
    """The autograd wrappers: window form and explicit patch-matrix form of conv2 (with the fused BN1 + clip) agree to
    bf16 rounding of the operands (both round the same values; only summation order differs)."""
    ops = _ops()
    rng = np.random.default_rng(5)
    B, Ti, Fi, C = 2, 41, 79, 32
    res = {}
    for use in (True, False):
        ops.USE_CONV_WINDOW = use
        x = torch.tensor(rng.standard_normal((B, Ti, Fi, C)) * 4 if use else res["x"], dtype=torch.float32, device="cuda").requires_grad_(True)
        if use:
            res["x"] = x.detach().cpu().numpy()
            res["w"] = rng.standard_normal((10, 5, C, C)) * 0.05
            res["b"] = rng.standard_normal(C) * 0.1
            res["g"] = rng.uniform(0.6, 1.4, C)
            res["be"] = rng.standard_normal(C) * 0.3
            res["dy"] = None
        mk = lambda k: torch.tensor(res[k], dtype=torch.float32, device="cuda").requires_grad_(True)
        w, b, g, be = mk("w"), mk("b"), mk("g"), mk("be")
        y = ops.conv2d_gemm(x, w, b, 2, 1, g, be, 1e-3, 20.0)
        if res["dy"] is None:
            res["dy"] = rng.standard_normal(tuple(y.shape))
        (y * torch.tensor(res["dy"], dtype=torch.float32, device="cuda")).sum().backward()
        torch.cuda.synchronize()
        res[use] = [t.detach().cpu().numpy() for t in (y, x.grad, w.grad, b.grad, g.grad, be.grad)]
    ops.USE_CONV_WINDOW = True
    for got, ref, name in zip(res[True], res[False], ("y", "dx", "dw", "db", "dgamma", "dbeta")):
        assert relerr(got, ref) < 2e-2, name


@pytest.mark.parametrize("order", ["nhwc", "cf", "fc"])
def test_bn_clip_fwd_bwd_vs_oracle(order):
    ops = _ops()
    rng = np.random.default_rng(12)
    B, T, Fq, C = 3, 7, 5, 4
    x = torch.tensor(rng.standard_normal((B, T, Fq, C)) * 8 + 3, dtype=torch.float32, device="cuda", requires_grad=True)
    gm = torch.tensor(rng.uniform(0.5, 1.5, C), dtype=torch.float32, device="cuda", requires_grad=True)
    bt = torch.tensor(rng.standard_normal(C), dtype=torch.float32, device="cuda", requires_grad=True)
    if order == "nhwc":
        y = ops.bn_clip_nhwc(x, gm, bt, 1e-3, 20.0)
    else:
        y = ops.bn_clip_to_rnn_layout(x, gm, bt, 1e-3, 20.0, order, torch.float32)
    g = torch.tensor(rng.standard_normal(tuple(y.shape)), dtype=torch.float32, device="cuda")
    (y * g).sum().backward()
    xr = x.detach().double().cpu().requires_grad_(True)
    gr = gm.detach().double().cpu().requires_grad_(True)
    br = bt.detach().double().cpu().requires_grad_(True)
    yr = O.relux(O.batch_norm_train(xr, gr, br, 1e-3, 3), 20.0)
    if order == "cf":
        yr = yr.permute(1, 0, 3, 2).reshape(T, B, C * Fq)
    elif order == "fc":
        yr = yr.permute(1, 0, 2, 3).reshape(T, B, Fq * C)
    (yr * g.double().cpu()).sum().backward()
    assert relerr(y.detach().cpu().numpy(), yr.detach().numpy()) < 1e-5
    assert relerr(x.grad.cpu().numpy(), xr.grad.numpy()) < 1e-3
    assert relerr(gm.grad.cpu().numpy(), gr.grad.numpy()) < 1e-3
    assert relerr(bt.grad.cpu().numpy(), br.grad.numpy()) < 1e-3


# ----------------------------------------------------------------------------------------- recurrent layers
def _layer_params(rng, H, I, G=1):
    mk = lambda *s: torch.tensor(rng.standard_normal(s) * (0.4 / np.sqrt(s[-1])), dtype=torch.float64)
    lp = {}
    for d in ("fw", "bw"):
        lp[d] = {"W": mk(G * H, I) * 3, "U": mk(G * H, H) * 2}
        if G == 1:
            lp[d]["B"] = torch.tensor(rng.standard_normal(H) * 0.1)
        else:
            lp[d]["BW"] = torch.tensor(rng.standard_normal(G * H) * 0.1)
            lp[d]["BU"] = torch.tensor(rng.standard_normal(G * H) * 0.1)
    return lp


@pytest.mark.parametrize("T,B,I,H,lens", [(5, 2, 6, 4, [5, 3]), (9, 3, 10, 16, [9, 9, 1]), (17, 4, 24, 40, [17, 2, 10, 16])])
def test_birnn_layer_fp32_vs_oracle(T, B, I, H, lens):
    ops = _ops()
    rng = np.random.default_rng(T * 7 + H)
    lp = _layer_params(rng, H, I)
    x64 = torch.tensor(rng.standard_normal((T, B, I)) * 2, requires_grad=True)
    for v in lp.values():
        for t in v.values():
            t.requires_grad_(True)
    ref = O.rnn_direction(x64, lens, lp["fw"], False) + O.rnn_direction(x64, lens, lp["bw"], True)
    g = torch.tensor(rng.standard_normal((T, B, H)))
    (ref * g).sum().backward()
    dev = "cuda"
    x = x64.detach().float().to(dev).requires_grad_(True)
    W = torch.cat([lp["fw"]["W"], lp["bw"]["W"]], 0).detach().float().to(dev).requires_grad_(True)
    U = torch.stack([lp["fw"]["U"], lp["bw"]["U"]], 0).detach().float().to(dev).requires_grad_(True)
    Bv = torch.stack([lp["fw"]["B"], lp["bw"]["B"]], 0).detach().float().to(dev).requires_grad_(True)
    lens_d = torch.tensor(lens, dtype=torch.int32, device=dev)
    out = ops.birnn_layer(x, W, U, Bv, lens_d, 0b10, False, torch.float32)
    (out * g.float().to(dev)).sum().backward()
    assert relerr(out.detach().cpu().numpy(), ref.detach().numpy()) < 1e-5
    assert relerr(x.grad.cpu().numpy(), x64.grad.numpy()) < 1e-3
    assert relerr(W.grad.cpu().numpy(), torch.cat([lp["fw"]["W"].grad, lp["bw"]["W"].grad], 0).numpy()) < 1e-3
    assert relerr(U.grad.cpu().numpy(), torch.stack([lp["fw"]["U"].grad, lp["bw"]["U"].grad], 0).numpy()) < 1e-3
    assert relerr(Bv.grad.cpu().numpy(), torch.stack([lp["fw"]["B"].grad, lp["bw"]["B"].grad], 0).numpy()) < 1e-3
    # masking: outputs are exactly zero past each utterance's length
    o = out.detach().cpu().numpy()
    for b, n in enumerate(lens):
        assert np.abs(o[n:, b]).max(initial=0.0) == 0.0


@pytest.mark.parametrize("T,B,I,H,lens", [(6, 2, 5, 3, [6, 4]), (11, 3, 12, 16, [11, 5, 1])])
def test_bigru_layer_fp32_vs_oracle(T, B, I, H, lens):
    ops = _ops()
    rng = np.random.default_rng(T * 5 + H)
    lp = _layer_params(rng, H, I, G=3)
    x64 = torch.tensor(rng.standard_normal((T, B, I)), requires_grad=True)
    for v in lp.values():
        for t in v.values():
            t.requires_grad_(True)
    ref = O.rnn_direction(x64, lens, lp["fw"], False, cell="gru") + O.rnn_direction(x64, lens, lp["bw"], True, cell="gru")
    g = torch.tensor(rng.standard_normal((T, B, H)))
    (ref * g).sum().backward()
    dev = "cuda"
    cat = lambda k: torch.cat([lp["fw"][k], lp["bw"][k]], 0).detach().float().to(dev).requires_grad_(True)
    stk = lambda k: torch.stack([lp["fw"][k], lp["bw"][k]], 0).detach().float().to(dev).requires_grad_(True)
    x = x64.detach().float().to(dev).requires_grad_(True)
    W, U, BW, BU = cat("W"), stk("U"), stk("BW"), stk("BU")
    out = ops.bigru_layer(x, W, U, BW, BU, torch.tensor(lens, dtype=torch.int32, device=dev), 0b10, False, torch.float32)
    (out * g.float().to(dev)).sum().backward()
    assert relerr(out.detach().cpu().numpy(), ref.detach().numpy()) < 1e-5
    assert relerr(x.grad.cpu().numpy(), x64.grad.numpy()) < 1e-3
    assert relerr(W.grad.cpu().numpy(), torch.cat([lp["fw"]["W"].grad, lp["bw"]["W"].grad], 0).numpy()) < 1e-3
    assert relerr(U.grad.cpu().numpy(), torch.stack([lp["fw"]["U"].grad, lp["bw"]["U"].grad], 0).numpy()) < 1e-3
    assert relerr(BW.grad.cpu().numpy(), torch.stack([lp["fw"]["BW"].grad, lp["bw"]["BW"].grad], 0).numpy()) < 1e-3
    assert relerr(BU.grad.cpu().numpy(), torch.stack([lp["fw"]["BU"].grad, lp["bw"]["BU"].grad], 0).numpy()) < 1e-3


def test_linear_and_errors():
    ops = _ops()
    rng = np.random.default_rng(2)
    x = torch.tensor(rng.standard_normal((50, 40)), dtype=torch.float32, device="cuda", requires_grad=True)
    W = torch.tensor(rng.standard_normal((29, 40)), dtype=torch.float32, device="cuda", requires_grad=True)
    b = torch.tensor(rng.standard_normal(29), dtype=torch.float32, device="cuda", requires_grad=True)
    y = ops.linear(x, W, b)
    y.square().sum().backward()
    xr, Wr, br = [t.detach().double().cpu().requires_grad_(True) for t in (x, W, b)]
    yr = xr @ Wr.t() + br
    yr.square().sum().backward()
    assert relerr(y.detach().cpu().numpy(), yr.detach().numpy()) < 1e-5
    for a, r in ((x, xr), (W, Wr), (b, br)):
        assert relerr(a.grad.cpu().numpy(), r.grad.numpy()) < 1e-4
    with pytest.raises(RuntimeError):
        ops.relux(torch.zeros(3))            # CPU tensor: no fallback


# ----------------------------------------------------------------------------------------- persistent tcgen05 recurrence
@pytest.mark.parametrize("T,B,I,H,ragged", [(6, 32, 64, 64, False), (9, 32, 128, 256, True), (5, 64, 72, 200, True),
                                             (4, 32, 256, 1760, False), (24, 32, 64, 800, True)])
def test_birnn_layer_bf16_persistent_vs_oracle(T, B, I, H, ragged):
    """bf16 operands / fp32 accumulation against the fp64 oracle run on the SAME bf16-rounded weights:
    <= 1e-2 of the output scale (north_star: bf16 recurrence <= 1e-2), gradients <= 3e-2."""
    ops = _ops()
    from deepspeech_b200._lib import lib
    assert lib.ds2_rnn_relu_persistent_supported(B, H, 2) == 1
    rng = np.random.default_rng(T * 13 + H)
    bf = lambda a: torch.tensor(a, dtype=torch.float32).to(torch.bfloat16).double()
    lp = {}
    for d in ("fw", "bw"):
        lp[d] = {"W": bf(rng.standard_normal((H, I)) * (1.0 / np.sqrt(I))).requires_grad_(True),
                 "U": bf(rng.standard_normal((H, H)) * (0.7 / np.sqrt(H))).requires_grad_(True),
                 "B": torch.tensor(rng.standard_normal(H) * 0.1).float().double().requires_grad_(True)}
    x64 = bf(rng.standard_normal((T, B, I))).requires_grad_(True)
    lens = [T] * B
    if ragged:
        lens = [T] + [int(v) for v in rng.integers(1, T + 1, size=B - 1)]
    ref = O.rnn_direction(x64, lens, lp["fw"], False) + O.rnn_direction(x64, lens, lp["bw"], True)
    g = torch.tensor(rng.standard_normal((T, B, H)))
    (ref * g).sum().backward()
    dev = "cuda"
    x = x64.detach().to(torch.bfloat16).to(dev).requires_grad_(True)
    W = torch.cat([lp["fw"]["W"], lp["bw"]["W"]], 0).detach().float().to(dev).requires_grad_(True)
    U = torch.stack([lp["fw"]["U"], lp["bw"]["U"]], 0).detach().float().to(dev).requires_grad_(True)
    Bv = torch.stack([lp["fw"]["B"], lp["bw"]["B"]], 0).detach().float().to(dev).requires_grad_(True)
    out = ops.birnn_layer(x, W, U, Bv, torch.tensor(lens, dtype=torch.int32, device=dev), 0b10, True, torch.float32)
    (out * g.float().to(dev)).sum().backward()
    torch.cuda.synchronize()
    o = out.detach().cpu().numpy()
    assert relerr(o, ref.detach().numpy()) < 1e-2
    for b, n in enumerate(lens):
        assert np.abs(o[n:, b]).max(initial=0.0) == 0.0
    # Gradients: a bf16 pre-activation that rounds across 0 (or 20) flips its ReLU mask, which changes single
    # gradient entries by 100%; element-wise parity with the fp64 oracle is therefore not defined for this path
    # (the fp32 path carries the <= 1e-3 gradient bar).  Check direction/scale here ...
    def cos(a, b):
        a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
        return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-300))
    assert cos(x.grad.float().cpu().numpy(), x64.grad.numpy()) > 0.99
    assert cos(W.grad.cpu().numpy(), torch.cat([lp["fw"]["W"].grad, lp["bw"]["W"].grad], 0).numpy()) > 0.99
    assert cos(U.grad.cpu().numpy(), torch.stack([lp["fw"]["U"].grad, lp["bw"]["U"].grad], 0).numpy()) > 0.99
    assert cos(Bv.grad.cpu().numpy(), torch.stack([lp["fw"]["B"].grad, lp["bw"]["B"].grad], 0).numpy()) > 0.99
    # ... and the backward recurrence kernel itself against the fp32 per-step kernels on the SAME stored
    # activations (identical masks): dz to bf16 rounding.
    from deepspeech_b200._lib import DS2_BF16, DS2_F32, check
    lens_d = torch.tensor(lens, dtype=torch.int32, device=dev)
    Ub = U.detach().to(torch.bfloat16)
    wx = ops.gemm_nt(x.detach().view(T * B, I), W.detach().to(torch.bfloat16))
    hs = torch.empty((2, T, B, H), dtype=torch.bfloat16, device=dev)
    wsb = lib.ds2_rnn_workspace_bytes(T, B, H, 2, DS2_BF16)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    p, st = ops._p, ops._stream
    check(lib.ds2_rnn_relu_fwd(p(wx), ops.SBN_SCALE, p(Ub), p(Bv.detach()), p(lens_d), p(hs), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "fwd")
    dout = torch.tensor(rng.standard_normal((T, B, H)), dtype=torch.float32, device=dev)
    dz_b = torch.empty((T, B, 2 * H), dtype=torch.bfloat16, device=dev)
    check(lib.ds2_rnn_relu_bwd(p(dout), p(Ub), p(lens_d), p(hs), p(dz_b), None, DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "bwd")
    hs32, U32 = hs.float(), Ub.float()
    dz_f = torch.empty((T, B, 2 * H), dtype=torch.float32, device=dev)
    check(lib.ds2_rnn_relu_bwd(p(dout), p(U32), p(lens_d), p(hs32), p(dz_f), None, DS2_F32, T, B, H, 2, 2, 20.0, None, 0, st()), "bwd32")
    torch.cuda.synchronize()
    assert relerr(dz_b.float().cpu().numpy(), dz_f.cpu().numpy()) < 2e-2


@pytest.mark.parametrize("T,B,I,H,ragged", [(5, 32, 64, 64, False), (9, 32, 128, 256, True), (7, 16, 72, 200, True),
                                             (4, 32, 256, 800, False), (20, 32, 64, 1024, True), (6, 64, 64, 328, True),
                                             (4, 32, 128, 1760, True), (3, 64, 64, 1096, True)])
def test_bigru_layer_bf16_persistent_vs_oracle(T, B, I, H, ragged):
    """Persistent tcgen05 GRU (bf16 operands, fp32 accumulation and gate math) against the fp64 oracle on the SAME
    bf16-rounded weights: outputs <= 1e-2 of the output scale; the GRU has no hard mask, so gradients are compared
    element-wise too (<= 3e-2 of each gradient's scale: activations and gate gradients are stored in bf16)."""
    ops = _ops()
    from deepspeech_b200._lib import lib
    assert lib.ds2_gru_persistent_supported(B, H, 2) == 1
    rng = np.random.default_rng(T * 17 + H)
    bf = lambda a: torch.tensor(a, dtype=torch.float32).to(torch.bfloat16).double()
    lp = {}
    for d in ("fw", "bw"):
        lp[d] = {"W": bf(rng.standard_normal((3 * H, I)) * (1.0 / np.sqrt(I))).requires_grad_(True),
                 "U": bf(rng.standard_normal((3 * H, H)) * (1.0 / np.sqrt(H))).requires_grad_(True),
                 "BW": torch.tensor(rng.standard_normal(3 * H) * 0.1).float().double().requires_grad_(True),
                 "BU": torch.tensor(rng.standard_normal(3 * H) * 0.1).float().double().requires_grad_(True)}
    x64 = bf(rng.standard_normal((T, B, I))).requires_grad_(True)
    lens = [T] * B
    if ragged:
        lens = [T] + [int(v) for v in rng.integers(1, T + 1, size=B - 1)]
    ref = O.rnn_direction(x64, lens, lp["fw"], False, cell="gru") + O.rnn_direction(x64, lens, lp["bw"], True, cell="gru")
    g = torch.tensor(rng.standard_normal((T, B, H)))
    (ref * g).sum().backward()
    dev = "cuda"
    x = x64.detach().to(torch.bfloat16).to(dev).requires_grad_(True)
    cat = lambda k: torch.cat([lp["fw"][k], lp["bw"][k]], 0).detach().float().to(dev).requires_grad_(True)
    stk = lambda k: torch.stack([lp["fw"][k], lp["bw"][k]], 0).detach().float().to(dev).requires_grad_(True)
    W, U, BW, BU = cat("W"), stk("U"), stk("BW"), stk("BU")
    n0 = lib.ds2_launch_count()
    out = ops.bigru_layer(x, W, U, BW, BU, torch.tensor(lens, dtype=torch.int32, device=dev), 0b10, True, torch.float32)
    (out * g.float().to(dev)).sum().backward()
    torch.cuda.synchronize()
    assert lib.ds2_launch_count() - n0 < 40          # one recurrence launch per pass, not one per step
    o = out.detach().cpu().numpy()
    assert relerr(o, ref.detach().numpy()) < 1e-2
    for b, n in enumerate(lens):
        assert np.abs(o[n:, b]).max(initial=0.0) == 0.0
    assert relerr(x.grad.float().cpu().numpy(), x64.grad.numpy()) < 3e-2
    assert relerr(W.grad.cpu().numpy(), torch.cat([lp["fw"]["W"].grad, lp["bw"]["W"].grad], 0).numpy()) < 3e-2
    assert relerr(U.grad.cpu().numpy(), torch.stack([lp["fw"]["U"].grad, lp["bw"]["U"].grad], 0).numpy()) < 3e-2
    assert relerr(BW.grad.cpu().numpy(), torch.stack([lp["fw"]["BW"].grad, lp["bw"]["BW"].grad], 0).numpy()) < 3e-2
    assert relerr(BU.grad.cpu().numpy(), torch.stack([lp["fw"]["BU"].grad, lp["bw"]["BU"].grad], 0).numpy()) < 3e-2


def test_rnn_persistent_refetch_path_and_v1_agree():
    """The round-2 exchange protocol of the persistent recurrence (ds2_rnn_persist2.cu): with DS2_RNN_EXP=1 the image
    check reports 'stale' once every 5th step, which drives the re-fetch path; the result must be bit-identical to the
    normal run.  The round-1 kernel (DS2_RNN_V1=1, fp32 slab exchange on both sides so that the arithmetic is the same)
    must agree with it bit for bit as well, forward and backward."""
    import os
    ops = _ops()
    from deepspeech_b200._lib import DS2_BF16, check, lib
    T, B, H = 23, 32, 1760
    g = torch.Generator(device="cuda").manual_seed(17)
    wx = torch.randn((T, B, 2 * H), device="cuda", generator=g) * 0.7
    U = (torch.randn((2, H, H), device="cuda", generator=g) * (0.6 / H ** 0.5)).to(torch.bfloat16)
    bias = torch.randn((2, H), device="cuda", generator=g) * 0.1
    dout = torch.randn((T, B, H), device="cuda", generator=g)
    lens = torch.tensor([T] + list(np.random.default_rng(3).integers(1, T + 1, size=B - 1)), dtype=torch.int32, device="cuda")
    p, st = ops._p, ops._stream
    wsb = lib.ds2_rnn_workspace_bytes(T, B, H, 2, DS2_BF16)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")

    def run():
        hs = torch.empty((2, T, B, H), dtype=torch.bfloat16, device="cuda")
        dz = torch.empty((T, B, 2 * H), dtype=torch.bfloat16, device="cuda")
        db = torch.empty((2, H), device="cuda")
        check(lib.ds2_rnn_relu_fwd(p(wx), 1.0, p(U), p(bias), p(lens), p(hs), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "fwd")
        check(lib.ds2_rnn_relu_bwd(p(dout), p(U), p(lens), p(hs), p(dz), p(db), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "bwd")
        torch.cuda.synchronize()
        return hs, dz, db

    saved = {k: os.environ.get(k) for k in ("DS2_RNN_EXP", "DS2_RNN_V1", "DS2_RNN_SLAB_F32")}
    try:
        os.environ["DS2_RNN_SLAB_F32"] = "1"
        lib.ds2_debug_reload_env()
        base = run()
        os.environ["DS2_RNN_EXP"] = "1"
        lib.ds2_debug_reload_env()
        forced = run()
        os.environ.pop("DS2_RNN_EXP")
        os.environ["DS2_RNN_V1"] = "1"
        lib.ds2_debug_reload_env()
        v1 = run()
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        lib.ds2_debug_reload_env()
    for a, b_, c in zip(base, forced, v1):
        assert torch.equal(a, b_)
        assert torch.equal(a[..., :] if a.dtype != torch.float32 else a, c) if a.dtype != torch.float32 else torch.allclose(a, c, rtol=1e-5, atol=1e-5)
