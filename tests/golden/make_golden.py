"""Generates tests/golden/*.npz.  The reference cannot be imported here (TensorFlow 1.x, python2 -
SURVEY F7/F8) and ships no golden vectors (F3), so these fixtures are produced by the CPU oracle and
cross-checked at generation time against an independent implementation (torch CPU fp64:
F.ctc_loss / conv2d / batch_norm).  They freeze the oracle's outputs so that a later edit of either the
oracle or the kernels cannot drift silently.  Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ds2_oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def ctc_fixture():
    rng = np.random.default_rng(2024)
    T, B, K = 24, 4, 29
    x = rng.standard_normal((T, B, K)).astype(np.float32)
    labels = [[3, 3, 7], [1], [], [5, 9, 5, 9, 2, 2]]
    lens = np.asarray([24, 10, 7, 20], np.int32)
    loss, grad = O.ctc_loss_and_grad_np(x, labels, lens)
    xt = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    tl = F.ctc_loss(torch.log_softmax(xt, 2), torch.tensor(sum(labels, []), dtype=torch.long), torch.tensor(lens, dtype=torch.long),
                    torch.tensor([len(l) for l in labels]), blank=28, reduction="none")
    tl.sum().backward()
    assert np.allclose(loss, tl.detach().numpy(), rtol=1e-10) and np.allclose(grad, xt.grad.numpy(), atol=1e-10)
    Lmax = max(len(l) for l in labels)
    lab = np.zeros((B, Lmax), np.int32)
    for b, l in enumerate(labels):
        lab[b, :len(l)] = l
    np.savez(os.path.join(HERE, "ctc_small.npz"), logits=x, labels=lab, label_lens=np.asarray([len(l) for l in labels], np.int32),
             seq_lens=lens, loss=loss, grad=grad)


def model_fixture():
    rng = np.random.default_rng(77)
    nf, H, L, B, T = 2, 8, 2, 2, 57
    p, layers = O.init_params(rng, num_filters=nf, num_hidden=H, num_layers=L)
    feats = rng.standard_normal((B, T, 161)).astype(np.float32)
    seq_lens = np.asarray([57, 49], np.int32)
    labels = [[4, 11], [20]]
    leaves = {}
    for k, v in p.items():
        leaves[k] = v.requires_grad_(True)
    for l, lp in enumerate(layers):
        for d, q in lp.items():
            for k, v in q.items():
                leaves["rnn.%d.%s.%s" % (l, d, k)] = v.requires_grad_(True)
    loss, logits = O.model_loss(torch.tensor(feats, dtype=torch.float64), seq_lens, labels, p, layers, dict(layout="NCHW", wiring="stacked"))
    loss.backward()
    out = {"feats": feats, "seq_lens": seq_lens, "labels": np.asarray([[4, 11], [20, 0]], np.int32), "label_lens": np.asarray([2, 1], np.int32),
           "loss": np.asarray(loss.item()), "logits": logits.detach().numpy()}
    for k, v in leaves.items():
        out["param:" + k] = v.detach().numpy()
        out["grad:" + k] = v.grad.numpy()
    np.savez(os.path.join(HERE, "model_small.npz"), **out)


def gru_fixture():
    """One bidirectional GRU layer (build-defined cell), cross-checked against torch.nn.GRU: with
    seq-BN as written the input projection is scaled by s = 1/sqrt(1+1e-5), i.e. nn.GRU with weight_ih = s*W."""
    rng = np.random.default_rng(31)
    T, B, I, H = 9, 3, 5, 4
    lp = {}
    for d in ("fw", "bw"):
        lp[d] = {"W": torch.tensor(rng.standard_normal((3 * H, I)) * 0.5), "U": torch.tensor(rng.standard_normal((3 * H, H)) * 0.5),
                 "BW": torch.tensor(rng.standard_normal(3 * H) * 0.1), "BU": torch.tensor(rng.standard_normal(3 * H) * 0.1)}
        for v in lp[d].values():
            v.requires_grad_(True)
    x = torch.tensor(rng.standard_normal((T, B, I)), requires_grad=True)
    lens = [T, T, T]                                    # full lengths for the nn.GRU cross-check
    out = O.rnn_direction(x, lens, lp["fw"], False, cell="gru") + O.rnn_direction(x, lens, lp["bw"], True, cell="gru")
    g = torch.nn.GRU(I, H, bidirectional=True).double()
    s = 1.0 / np.sqrt(1.0 + 1e-5)
    with torch.no_grad():
        for d, suf in (("fw", ""), ("bw", "_reverse")):
            getattr(g, "weight_ih_l0" + suf).copy_(lp[d]["W"] * s)
            getattr(g, "weight_hh_l0" + suf).copy_(lp[d]["U"])
            getattr(g, "bias_ih_l0" + suf).copy_(lp[d]["BW"])
            getattr(g, "bias_hh_l0" + suf).copy_(lp[d]["BU"])
    ref, _ = g(x.detach())
    assert np.allclose(out.detach().numpy(), (ref[:, :, :H] + ref[:, :, H:]).detach().numpy(), atol=1e-12)
    lens = [9, 6, 2]                                    # the fixture itself is ragged
    out = O.rnn_direction(x, lens, lp["fw"], False, cell="gru") + O.rnn_direction(x, lens, lp["bw"], True, cell="gru")
    gout = torch.tensor(rng.standard_normal((T, B, H)))
    (out * gout).sum().backward()
    z = {"x": x.detach().numpy(), "lens": np.asarray(lens, np.int32), "gout": gout.numpy(), "out": out.detach().numpy(), "dx": x.grad.numpy()}
    for d in ("fw", "bw"):
        for k, v in lp[d].items():
            z["%s.%s" % (d, k)] = v.detach().numpy()
            z["grad:%s.%s" % (d, k)] = v.grad.numpy()
    np.savez(os.path.join(HERE, "gru_small.npz"), **z)


def optim_fixture():
    """Three Adam + EMA steps with a decaying learning rate (oracle restatement of src/deepSpeech_train.py:231-250,
    :430, :457-465), cross-checked against torch.optim.Adam in the eps -> 0 limit at generation time."""
    rng = np.random.default_rng(8)
    w0 = rng.standard_normal(37)
    grads = rng.standard_normal((3, 37))
    w, m, v, sh = w0.copy(), np.zeros(37), np.zeros(37), w0.copy()
    wt, mt, vt = w0.copy(), np.zeros(37), np.zeros(37)
    p = torch.tensor(w0, requires_grad=True)
    opt = torch.optim.Adam([p], lr=1e-2, eps=1e-30)
    ws, shs, lrs = [], [], []
    for step in range(1, 4):
        lr = O.decayed_learning_rate(1e-2, step - 1, 2, 0.5)
        w, m, v, sh = O.adam_ema_step_np(w, grads[step - 1], m, v, sh, step, lr, 0.9)
        ws.append(w.copy()); shs.append(sh.copy()); lrs.append(lr)
        wt, mt, vt, _ = O.adam_ema_step_np(wt, grads[step - 1], mt, vt, None, step, 1e-2, 0.9, eps=1e-30)
        p.grad = torch.tensor(grads[step - 1])
        opt.step()
        assert np.abs(wt - p.detach().numpy()).max() < 1e-12
    np.savez(os.path.join(HERE, "optim_small.npz"), w0=w0, grads=grads, w=np.stack(ws), shadow=np.stack(shs), lr=np.asarray(lrs))


def decode_fixture():
    rng = np.random.default_rng(12)
    T, B, K = 40, 5, 29
    x = rng.standard_normal((T, B, K)).astype(np.float32)
    x[:, :, K - 1] += 1.5
    for b in range(B):
        for t in range(1, T):
            if rng.random() < 0.35:
                x[t, b] = x[t - 1, b]
    lens = np.asarray([40, 31, 1, 0, 17], np.int32)
    dec = O.ctc_greedy_decode_np(x, lens)
    labels = [[int(c) for c in rng.integers(0, 28, size=n)] for n in (12, 5, 3, 2, 9)]
    cer = O.char_error_rates(labels, dec)
    pack = lambda seqs: (np.asarray([s + [0] * (max(1, max(map(len, seqs))) - len(s)) for s in seqs], np.int32), np.asarray([len(s) for s in seqs], np.int32))
    d, dl = pack(dec)
    lab, ll = pack(labels)
    np.savez(os.path.join(HERE, "decode_small.npz"), logits=x, seq_lens=lens, decoded=d, decoded_lens=dl, labels=lab, label_lens=ll,
             cer=np.asarray(cer))


if __name__ == "__main__":
    ctc_fixture()
    model_fixture()
    gru_fixture()
    optim_fixture()
    decode_fixture()
    print("golden fixtures written to", HERE)
