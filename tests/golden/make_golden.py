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


if __name__ == "__main__":
    ctc_fixture()
    model_fixture()
    print("golden fixtures written to", HERE)
