"""GPU parity tests, model level: the drop-in surface (inference / loss / stacked_brnn / cells / BN helpers)
against the CPU oracle on identical synthetic inputs and weights.  Tolerances per BASELINE.json:
CTC loss <= 1e-4 relative (fp32), gradients <= 1e-3 relative, bf16 path <= 1e-2 on logits."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import ds2_oracle as O

pytestmark = pytest.mark.gpu


def relerr(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _oracle_run(feats, seq_lens, sparse, p, layers, cfg):
    for t in p.values():
        t.requires_grad_(True)
    for lp in layers:
        for d in lp.values():
            for t in d.values():
                t.requires_grad_(True)
    seqs = O.sparse_to_label_seqs(*sparse)
    loss, logits = O.model_loss(torch.tensor(feats, dtype=torch.float64), seq_lens, seqs, p, layers, cfg)
    loss.backward()
    return loss.item(), logits.detach().numpy()


def _small_case(seed, B=3, T=61, nf=4, H=16, L=2, cell="relu_clip", wiring="stacked", ragged=True):
    rng = np.random.default_rng(seed)
    p, layers = O.init_params(rng, num_filters=nf, num_hidden=H, num_layers=L, cell=cell, wiring=wiring)
    # non-trivial BN affine + biases so their gradients are exercised
    for k in ("conv1.gamma", "conv2.gamma"):
        p[k] = torch.tensor(rng.uniform(0.6, 1.4, nf))
    for k in ("conv1.beta", "conv2.beta"):
        p[k] = torch.tensor(rng.normal(size=nf) * 0.3)
    for lp in layers:
        for d in lp.values():
            for k in d:
                if k.startswith("B"):
                    d[k] = torch.tensor(rng.normal(size=tuple(d[k].shape)) * 0.1)
    feats = rng.standard_normal((B, T, 161)).astype(np.float32)
    seq_lens = np.asarray([T] + [int(v) for v in rng.integers(45, T + 1, size=B - 1)] if ragged else [T] * B, np.int32)
    dense = np.zeros((B, 4), np.int64)
    for b in range(B):
        n = int(rng.integers(1, 4))
        dense[b, :n] = rng.integers(1, 28, size=n)
    idx = np.argwhere(dense != 0)
    vals = dense[dense != 0].astype(np.int32)
    return p, layers, feats, seq_lens, (idx, vals, np.asarray(dense.shape))


@pytest.mark.parametrize("layout,wiring,cell", [("NCHW", "stacked", "relu_clip"), ("NCHW", "as_written", "relu_clip"),
                                                ("NHWC", "two_stack", "relu_clip"), ("NCHW", "stacked", "gru")])
def test_model_fp32_loss_and_grads_vs_oracle(layout, wiring, cell):
    from deepspeech_b200 import deepSpeech, deepSpeech_NCHW
    mod = deepSpeech_NCHW if layout == "NCHW" else deepSpeech
    p, layers, feats, seq_lens, sparse = _small_case(17, cell=cell, wiring=wiring)
    ref_loss, ref_logits = _oracle_run(feats, seq_lens, sparse, p, layers, dict(layout=layout, wiring=wiring, cell=cell))
    params = SimpleNamespace(use_fp16=False, num_filters=4, batch_size=3, num_hidden=16, num_rnn_layers=2, engine="mkl",
                             rnn_type="bi-dir", wiring=wiring, cell=cell, seed=99)
    m = mod.model(params)
    m.load_oracle_params({k: v.detach() for k, v in p.items()},
                         [{d: {k: v.detach() for k, v in q.items()} for d, q in lp.items()} for lp in layers])
    m.zero_grad(set_to_none=True)
    logits = mod.inference(None, torch.tensor(feats, device="cuda"), seq_lens, params)
    loss = mod.loss(logits, sparse, seq_lens)
    loss.backward()
    torch.cuda.synchronize()
    assert abs(loss.item() - ref_loss) <= 1e-4 * abs(ref_loss)
    assert relerr(logits.detach().cpu().numpy(), ref_logits) < 1e-4
    g = m.oracle_grads()
    for k in ("conv1.weights", "conv1.gamma", "conv1.beta", "conv2.weights", "conv2.gamma", "conv2.beta",
              "softmax_linear.weights", "softmax_linear.biases"):
        assert relerr(g[k], p[k].grad.numpy().reshape(g[k].shape)) < 1e-3, k
    # conv biases feed a batch-stat BN: their true gradient is 0; compare absolutely
    for k in ("conv1.biases", "conv2.biases"):
        assert np.abs(g[k]).max() < 1e-4 * max(1.0, np.abs(g["conv2.weights"]).max()), k
    live = [len(layers) - 1] if wiring == "as_written" else range(len(layers))
    for l in live:
        lp = layers[l]
        assert relerr(g["rnn.%d.W" % l], torch.cat([lp["fw"]["W"].grad, lp["bw"]["W"].grad], 0).numpy()) < 1e-3
        assert relerr(g["rnn.%d.U" % l], torch.stack([lp["fw"]["U"].grad, lp["bw"]["U"].grad], 0).numpy()) < 1e-3
        bk = "BW" if cell == "gru" else "B"
        assert relerr(g["rnn.%d.B" % l], torch.stack([lp["fw"][bk].grad, lp["bw"][bk].grad], 0).numpy()) < 1e-3
    if wiring == "as_written":
        assert g["rnn.0.W"] is None           # dead layers get no gradient, as in the reference (TF prunes them)


def test_model_uni_dir_vs_oracle():
    """rnn_type='uni-dir' (src/deepSpeech.py:173-177): the forward stack alone; the backward-direction weights get
    no gradient."""
    from deepspeech_b200 import deepSpeech as mod
    p, layers, feats, seq_lens, sparse = _small_case(29, wiring="two_stack")
    cfg = dict(layout="NHWC", wiring="two_stack", bidirectional=False)
    ref_loss, ref_logits = _oracle_run(feats, seq_lens, sparse, p, layers, cfg)
    params = SimpleNamespace(use_fp16=False, num_filters=4, batch_size=3, num_hidden=16, num_rnn_layers=2, engine="mkl",
                             rnn_type="uni-dir", wiring="two_stack", seed=3)
    m = mod.model(params)
    m.load_oracle_params({k: v.detach() for k, v in p.items()},
                         [{d: {k: v.detach() for k, v in q.items()} for d, q in lp.items()} for lp in layers])
    m.zero_grad(set_to_none=True)
    logits = mod.inference(None, torch.tensor(feats, device="cuda"), seq_lens, params)
    loss = mod.loss(logits, sparse, seq_lens)
    loss.backward()
    torch.cuda.synchronize()
    assert abs(loss.item() - ref_loss) <= 1e-4 * abs(ref_loss)
    assert relerr(logits.detach().cpu().numpy(), ref_logits) < 1e-4
    g = m.oracle_grads()
    H = 16
    for l, lp in enumerate(layers):
        assert relerr(g["rnn.%d.W" % l][:H], lp["fw"]["W"].grad.numpy()) < 1e-3
        assert relerr(g["rnn.%d.U" % l][0], lp["fw"]["U"].grad.numpy()) < 1e-3
        assert lp["bw"]["U"].grad is None or float(lp["bw"]["U"].grad.abs().max()) == 0.0
        assert np.abs(g["rnn.%d.U" % l][1]).max() == 0.0


@pytest.mark.parametrize("cell", ["relu_clip", "gru"])
def test_model_bf16_persistent_kernels_vs_oracle(cell):
    """bf16 model at a batch the persistent recurrence kernels take (B=16): logits <= 1e-2 against the fp64 oracle,
    and the per-layer recurrence really is one launch (launch count far below T per layer)."""
    from deepspeech_b200 import deepSpeech_NCHW as mod
    from deepspeech_b200._lib import lib
    p, layers, feats, seq_lens, sparse = _small_case(31, B=16, T=101, nf=8, H=64, L=2, cell=cell, ragged=True)
    ref_loss, ref_logits = _oracle_run(feats, seq_lens, sparse, p, layers, dict(layout="NCHW", wiring="stacked", cell=cell))
    params = SimpleNamespace(use_fp16=True, num_filters=8, batch_size=16, num_hidden=64, num_rnn_layers=2, engine="mkl",
                             rnn_type="bi-dir", wiring="stacked", cell=cell, seed=5)
    m = mod.model(params)
    m.load_oracle_params({k: v.detach() for k, v in p.items()},
                         [{d: {k: v.detach() for k, v in q.items()} for d, q in lp.items()} for lp in layers])
    m.zero_grad(set_to_none=True)
    n0 = lib.ds2_launch_count()
    logits = mod.inference(None, torch.tensor(feats, device="cuda"), seq_lens, params)
    loss = mod.loss(logits, sparse, seq_lens)
    loss.backward()
    torch.cuda.synchronize()
    assert lib.ds2_launch_count() - n0 < 150
    assert relerr(logits.detach().cpu().numpy(), ref_logits) < 1e-2
    assert abs(loss.item() - ref_loss) <= 2e-2 * abs(ref_loss)
    g = m.oracle_grads()
    lp = layers[0]
    a = g["rnn.0.U"].ravel()
    b = torch.stack([lp["fw"]["U"].grad, lp["bw"]["U"].grad], 0).numpy().ravel()
    assert float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b))) > 0.99


def test_model_bf16_logits_vs_fp32_oracle():
    """bf16 tensor-core path against the fp64 oracle: <= 1e-2 on logits (of the logits' scale)."""
    from deepspeech_b200 import deepSpeech_NCHW as mod
    p, layers, feats, seq_lens, sparse = _small_case(23, B=8, T=81, nf=4, H=64, L=2, ragged=True)
    ref_loss, ref_logits = _oracle_run(feats, seq_lens, sparse, p, layers, dict(layout="NCHW", wiring="stacked"))
    params = SimpleNamespace(use_fp16=True, num_filters=4, batch_size=8, num_hidden=64, num_rnn_layers=2, engine="mkl",
                             rnn_type="bi-dir", wiring="stacked", seed=5)
    m = mod.model(params)
    m.load_oracle_params({k: v.detach() for k, v in p.items()},
                         [{d: {k: v.detach() for k, v in q.items()} for d, q in lp.items()} for lp in layers])
    m.zero_grad(set_to_none=True)
    logits = mod.inference(None, torch.tensor(feats, device="cuda"), seq_lens, params)
    loss = mod.loss(logits, sparse, seq_lens)
    loss.backward()
    torch.cuda.synchronize()
    assert relerr(logits.detach().cpu().numpy(), ref_logits) < 1e-2
    assert abs(loss.item() - ref_loss) <= 2e-2 * abs(ref_loss)
    g = m.oracle_grads()
    lp = layers[0]
    assert relerr(g["rnn.0.U"], torch.stack([lp["fw"]["U"].grad, lp["bw"]["U"].grad], 0).numpy()) < 5e-2


def test_surface_ops_vs_oracle():
    from deepspeech_b200 import custom_ops as C
    C.reset_variable_store()
    rng = np.random.default_rng(3)
    x = torch.tensor(rng.standard_normal((2, 4, 9, 5)) * 15, dtype=torch.float32, device="cuda")
    assert torch.equal(C.relux(x, capping=20).cpu(), x.cpu().clamp(0, 20))
    assert torch.equal(C.relux(x).cpu(), x.cpu().clamp(min=0))
    y = C.batch_norm2(x, data_format="NCHW", scope="t1")
    ref = O.batch_norm_train(x.double().cpu(), torch.ones(4, dtype=torch.float64), torch.zeros(4, dtype=torch.float64), 1e-3, 1)
    assert relerr(y.detach().cpu().numpy(), ref.numpy()) < 1e-5
    y = C.batch_norm(x, scope="t2")
    ref = O.batch_norm_train(x.double().cpu(), torch.ones(5, dtype=torch.float64), torch.zeros(5, dtype=torch.float64), 1e-5, 3)
    assert relerr(y.detach().cpu().numpy(), ref.numpy()) < 1e-5
    z = torch.tensor(rng.standard_normal((6, 7)), dtype=torch.float32, device="cuda")
    assert relerr(C.seq_batch_norm(z).cpu().numpy(), O.seq_batch_norm_as_written(z.double().cpu()).numpy()) < 1e-6
    with pytest.raises(ValueError):
        C._linear(None, 4, True)
    with pytest.raises(ValueError):
        C._linear([torch.zeros(2, 3, 4, device="cuda")], 4, True)
    # single-step cell against the oracle's cell
    cell = C.CustomRNNCell2(8)
    assert cell.output_size == 8 and cell.state_size == 8
    xin = torch.tensor(rng.standard_normal((3, 6)), dtype=torch.float32, device="cuda")
    h0 = torch.tensor(rng.uniform(0, 2, (3, 8)), dtype=torch.float32, device="cuda")
    out, st = cell(xin, h0, scope="cellA")
    W, U, Bv = cell.variables("cellA", 6)
    ref = O.rnn_cell_step(xin.double().cpu(), h0.double().cpu(), W.detach().double().cpu(), U.detach().double().cpu(),
                          Bv.detach().double().cpu())
    assert out is st and relerr(out.detach().cpu().numpy(), ref.numpy()) < 1e-5
    assert cell.zero_state(3).shape == (3, 8)
    # stacked_brnn: as_written == last layer only, on the original inputs
    T, B, I, H, L = 7, 2, 6, 8, 3
    cells = [C.CustomRNNCell2(H)] * L
    xs = torch.tensor(rng.standard_normal((T, B, I)), dtype=torch.float32, device="cuda")
    lens = np.asarray([7, 4], np.int32)
    out = C.stacked_brnn(cells, cells, H, L, xs, lens, B, wiring="as_written", scope="sb")
    last = {d: dict(zip("WUB", [v.detach().double().cpu() for v in cells[0].variables("sb/brnn-2/bidirectional_rnn/%s" % d, I)]))
            for d in ("fw", "bw")}
    ref = O.stacked_brnn(xs.double().cpu(), lens, [last], "as_written")
    assert relerr(out.detach().cpu().numpy(), ref.numpy()) < 1e-5
    with pytest.raises(ValueError):
        C.stacked_brnn(cells, cells, H, L, xs, lens, B + 1)
    # MkldnnRNNCell: API shape (flat blob, (output, output_h))
    from deepspeech_b200.mkldnn_rnn_op import MkldnnRNNCell
    mc = MkldnnRNNCell(None, 8, input_size=6)
    assert mc.model.params_size() == 8 * 6 + 64 + 16
    o, oh = mc(xin, h0, scope="mk")
    ref = torch.relu((xin.double().cpu().sum(1, keepdim=True) + h0.double().cpu().sum(1, keepdim=True) + 2) / mc.param_size)
    assert relerr(o.detach().cpu().numpy(), ref.expand(3, 8).numpy()) < 1e-5


def test_dummy_step_runs_like_the_reference_loop():
    """One dummy=1 training step (src/deepSpeech_train.py:305-325): inputs -> inference -> loss -> grads,
    loss finite, at a reduced model size; labels go through the Q8 transform."""
    from deepspeech_b200 import deepSpeech_NCHW as mod
    from deepspeech_b200.deepSpeech_dummy import DummyData
    params = SimpleNamespace(use_fp16=False, num_filters=4, batch_size=2, num_hidden=16, num_rnn_layers=2, engine="mkl",
                             rnn_type="bi-dir", wiring="stacked", seed=7)
    data = DummyData(2, seed=1)
    feats, idx, vals, shape, seq_lens = data.inputs()
    logits = mod.inference(None, torch.from_numpy(feats).cuda(), seq_lens, params)
    assert logits.shape == (16, 2, 29)
    loss = mod.loss(logits, (idx, vals, shape), seq_lens)
    loss.backward()
    assert np.isfinite(loss.item())
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in mod.model(params).parameters())
