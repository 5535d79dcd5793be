"""Pins for the CPU oracle (oracle/ds2_oracle.py): closed-form known answers and an independent
implementation (torch CPU fp64).  The reference ships no tests or golden vectors,
so these pins are what stands between the oracle and 'unpinned'."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import ds2_oracle as O


def test_get_rnn_seqlen_matches_valid_conv_sizes():
    for T in O.UTT_LENGTHS + [39, 40, 41, 57, 58, 123, 6034]:
        assert O.get_rnn_seqlen([T])[0] == O.conv_out_lens(T)[1]


def test_ctc_closed_forms():
    rng = np.random.default_rng(0)
    # (i) T=1, L=1: loss = -log softmax(x)[l]
    x = rng.standard_normal((1, 1, 29))
    loss, grad = O.ctc_loss_and_grad_np(x, [[5]], [1])
    ls = x[0, 0] - np.log(np.exp(x[0, 0]).sum())
    assert abs(loss[0] + ls[5]) < 1e-12
    # (ii) uniform logits, no repeats: p = C(T+L, 2L) / K^T
    for T, L in [(1, 1), (2, 1), (2, 2), (3, 2), (7, 3), (12, 5)]:
        lab = list(range(1, L + 1))
        loss, grad = O.ctc_loss_and_grad_np(np.zeros((T, 1, 29)), [lab], [T])
        assert abs(loss[0] - (T * math.log(29) - math.log(math.comb(T + L, 2 * L)))) < 1e-9
    # (iii) empty label: loss = -sum_t log y_t[blank]
    x = rng.standard_normal((6, 1, 29))
    loss, grad = O.ctc_loss_and_grad_np(x, [[]], [6])
    ly = x[:, 0] - np.log(np.exp(x[:, 0]).sum(1, keepdims=True))
    assert abs(loss[0] + ly[:, 28].sum()) < 1e-12
    # (iv) grad rows sum to 0 inside, are 0 outside
    x = rng.standard_normal((9, 2, 29))
    loss, grad = O.ctc_loss_and_grad_np(x, [[3, 3, 4], [1]], [9, 5])
    assert np.abs(grad.sum(2)).max() < 1e-12
    assert np.abs(grad[5:, 1]).max() == 0.0


@pytest.mark.parametrize("T,B,Lmax", [(20, 4, 6), (50, 3, 20), (64, 5, 1)])
def test_ctc_vs_torch_independent(T, B, Lmax):
    rng = np.random.default_rng(T)
    x = rng.standard_normal((T, B, 29))
    lens = rng.integers(T // 2 + Lmax, T + 1, size=B)
    labels = [list(rng.integers(0, 28, size=rng.integers(1, Lmax + 1))) for _ in range(B)]
    loss, grad = O.ctc_loss_and_grad_np(x, labels, lens)
    xt = torch.tensor(x, requires_grad=True)
    tl = F.ctc_loss(torch.log_softmax(xt, 2), torch.tensor(sum(labels, [])), torch.tensor(lens),
                    torch.tensor([len(l) for l in labels]), blank=28, reduction="none")
    tl.sum().backward()
    assert np.allclose(loss, tl.detach().numpy(), rtol=1e-10, atol=1e-10)
    assert np.allclose(grad, xt.grad.numpy(), rtol=1e-8, atol=1e-10)
    # autograd-able restatement agrees with the analytic one
    xt2 = torch.tensor(x, requires_grad=True)
    l2 = O.ctc_loss_torch(xt2, labels, lens)
    l2.sum().backward()
    assert np.allclose(loss, l2.detach().numpy(), rtol=1e-10)
    assert np.allclose(grad, xt2.grad.numpy(), rtol=1e-8, atol=1e-10)


def test_ctc_too_short_raises():
    with pytest.raises(ValueError):
        O.ctc_loss_and_grad_np(np.zeros((3, 1, 29)), [[1, 1, 2]], [3])   # needs 4 frames


def test_label_transform_q8():
    # zeros already dropped by the sparse conversion; repeats collapsed; >=28 terminates
    idx = [[0, 0], [0, 1], [0, 3], [0, 4], [0, 5], [1, 0], [1, 2]]
    vals = [5, 5, 7, 28, 28, 9, 28]
    seqs = O.sparse_to_label_seqs(idx, vals, [2, 6])
    assert seqs == [[5, 7], [9]]
    assert O.sparse_to_label_seqs(idx, vals, [2, 6], preprocess_collapse_repeated=False) == [[5, 5, 7], [9]]
    with pytest.raises(ValueError):
        O.sparse_to_label_seqs([[0, 0], [0, 1]], [28, 3], [1, 2])


def test_dummy_batch_shapes():
    rng = np.random.default_rng(1234)
    feats, idx, vals, shape, lens = O.dummy_batch(rng, 4, 1)
    assert feats.shape == (4, 200, 161) and feats.dtype == np.float32
    assert shape.tolist() == [4, 17] and (lens == 200).all()
    seqs = O.sparse_to_label_seqs(idx, vals, shape)
    assert all(s == seqs[0] for s in seqs) and all(0 < v < 28 for v in seqs[0])


def test_frontend_vs_independent_nchw_and_nhwc():
    rng = np.random.default_rng(3)
    p, _ = O.init_params(rng, num_filters=4, num_hidden=8, num_layers=1)
    p["conv1.gamma"] = torch.tensor(rng.uniform(0.5, 1.5, 4)); p["conv1.beta"] = torch.tensor(rng.normal(size=4))
    feats = torch.tensor(rng.standard_normal((3, 61, 161)))
    out, im = O.conv_frontend(feats, p, "NCHW", return_intermediates=True)
    T1, T2 = O.conv_out_lens(61)
    assert im["y1"].shape == (3, 4, T1, 79) and im["y2"].shape == (3, 4, T2, 75) and out.shape == (T2, 3, 300)
    # independent: explicit correlation loop for one output element + F.batch_norm
    w1 = p["conv1.weights"].numpy()
    b, c, t, f = 1, 2, 5, 7
    ref = (feats.numpy()[b, 2 * t:2 * t + 20, 2 * f:2 * f + 5] * w1[:, :, 0, c]).sum() - 0.05
    assert abs(im["y1"][b, c, t, f].item() - ref) < 1e-12
    bn = F.batch_norm(im["y1"], None, None, p["conv1.gamma"], p["conv1.beta"], training=True, eps=1e-3)
    assert torch.allclose(im["a1"], bn.clamp(0, 20), atol=1e-12)
    # flatten order: NCHW feature = c*75+f ; NHWC feature = f*nf+c
    out2 = O.conv_frontend(feats, p, "NHWC")
    a2 = im["a2"]
    assert out[3, 1, 2 * 75 + 11].item() == a2[1, 2, 3, 11].item()
    assert out2.shape == out.shape


def test_rnn_direction_masking_and_reversal():
    rng = np.random.default_rng(5)
    T, B, I, H = 7, 3, 6, 4
    x = torch.tensor(rng.standard_normal((T, B, I)))
    cp = {"W": torch.tensor(rng.normal(size=(H, I))), "U": torch.tensor(rng.normal(size=(H, H)) * 0.3),
          "B": torch.tensor(rng.normal(size=H))}
    lens = [7, 4, 1]
    fw = O.rnn_direction(x, lens, cp, False)
    bw = O.rnn_direction(x, lens, cp, True)
    for b, n in enumerate(lens):
        assert fw[n:, b].abs().max().item() == 0.0 if n < T else True
        # bw over the valid prefix == fw over the flipped prefix, flipped back
        xb = x[:n, b:b + 1].flip(0)
        ref = O.rnn_direction(xb, [n], cp, False).flip(0)
        assert torch.allclose(bw[:n, b:b + 1], ref, atol=1e-13)
        assert bw[n:, b].abs().sum().item() == 0.0
    # unclipped sanity against torch.nn.RNN(relu) on small activations (no seq-BN scale)
    rnn = torch.nn.RNN(I, H, nonlinearity="relu", bias=True).double()
    with torch.no_grad():
        rnn.weight_ih_l0.copy_(cp["W"] * 0.1); rnn.weight_hh_l0.copy_(cp["U"] * 0.1)
        rnn.bias_ih_l0.copy_(cp["B"] * 0.1); rnn.bias_hh_l0.zero_()
    cp2 = {k: v * 0.1 for k, v in cp.items()}
    mine = O.rnn_direction(x, [T] * B, cp2, False, sbn="none")
    assert torch.allclose(mine, rnn(x)[0], atol=1e-12)


def test_gru_vs_torch_gru():
    rng = np.random.default_rng(6)
    T, B, I, H = 6, 2, 5, 3
    gru = torch.nn.GRU(I, H).double()
    cp = {"W": gru.weight_ih_l0.detach(), "U": gru.weight_hh_l0.detach(),
          "BW": gru.bias_ih_l0.detach(), "BU": gru.bias_hh_l0.detach()}
    x = torch.tensor(rng.standard_normal((T, B, I)))
    mine = O.rnn_direction(x, [T] * B, cp, False, cell="gru", sbn="none")
    assert torch.allclose(mine, gru(x)[0], atol=1e-12)


def test_wirings():
    rng = np.random.default_rng(7)
    p, layers = O.init_params(rng, num_filters=1, num_hidden=5, num_layers=3, wiring="as_written")
    x = torch.tensor(rng.standard_normal((6, 2, 75)))
    lens = [6, 3]
    aw = O.stacked_brnn(x, lens, layers, "as_written")
    last = O.stacked_brnn(x, lens, layers[-1:], "stacked")
    assert torch.equal(aw, last)                       # Q1: only the last layer is live
    p, layers = O.init_params(rng, num_filters=1, num_hidden=5, num_layers=2, wiring="stacked")
    st = O.stacked_brnn(x, lens, layers, "stacked")
    ts = O.stacked_brnn(x, lens, layers, "two_stack")
    assert st.shape == ts.shape == (6, 2, 5) and not torch.allclose(st, ts)


def test_model_loss_backward_finite_difference():
    rng = np.random.default_rng(8)
    p, layers = O.init_params(rng, num_filters=2, num_hidden=6, num_layers=2)
    feats = torch.tensor(rng.standard_normal((2, 45, 161)))
    seq_lens = [45, 41]
    labels = [[3, 4], [7]]
    cfg = dict(layout="NCHW", wiring="stacked")
    U = layers[0]["fw"]["U"].requires_grad_(True)
    loss, _ = O.model_loss(feats, seq_lens, labels, p, layers, cfg)
    g, = torch.autograd.grad(loss, U)
    eps = 1e-6
    with torch.no_grad():
        U[1, 2] += eps
        lp, _ = O.model_loss(feats, seq_lens, labels, p, layers, cfg)
        U[1, 2] -= 2 * eps
        lm, _ = O.model_loss(feats, seq_lens, labels, p, layers, cfg)
        U[1, 2] += eps
    assert abs((lp - lm).item() / (2 * eps) - g[1, 2].item()) < 1e-6 * max(1.0, abs(g[1, 2].item()))
