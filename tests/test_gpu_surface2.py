"""Operator-surface rows that round 1 left partial: the whole-sequence op behind MkldnnRNNCell's flat
parameter blob (src/mkldnn_rnn_op.py:16-45, src/deepSpeech_NCHW.py:173-176) and the positive paths of `_linear` /
`CustomRNNCell` (src/custom_ops.py:14-33, :202-257).  GRU-free, fp32 path: <= 1e-4 against the fp64 oracle."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import ds2_oracle as O

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _blob(rng, H, I):
    W = rng.standard_normal((H, I)) / np.sqrt(I)
    U = rng.standard_normal((H, H)) * (0.5 / np.sqrt(H))
    bW, bU = rng.standard_normal(H) * 0.1, rng.standard_normal(H) * 0.1
    return np.concatenate([W.ravel(), U.ravel(), bW, bU]).astype(np.float32), W, U, bW + bU


@pytest.mark.parametrize("reverse", [False, True])
def test_mkldnn_sequence_op_matches_oracle_and_the_per_step_cell(reverse):
    from deepspeech_b200.mkldnn_rnn_op import MkldnnRNNCell
    rng = np.random.default_rng(3)
    T, B, I, H = 9, 5, 12, 16
    blob_np, W, U, b = _blob(rng, H, I)
    x_np = rng.standard_normal((T, B, I)).astype(np.float32)
    lens = [T, 3, 9, 1, 6]
    cell = MkldnnRNNCell(None, H, input_size=I)
    assert cell.param_size == H * I + H * H + 2 * H == blob_np.size
    blob = torch.tensor(blob_np, device="cuda", requires_grad=True)
    x = torch.tensor(x_np, device="cuda", requires_grad=True)
    out = cell.model.sequence(x, torch.tensor(lens, dtype=torch.int32, device="cuda"), blob, reverse, False)
    g = torch.tensor(rng.standard_normal((T, B, H)), dtype=torch.float32, device="cuda")
    (out * g).sum().backward()
    # oracle: plain ReLU, no seq-BN, no clip, dynamic_rnn masking
    cp = {"W": torch.tensor(W, requires_grad=True), "U": torch.tensor(U, requires_grad=True), "B": torch.tensor(b, requires_grad=True)}
    x64 = torch.tensor(x_np, dtype=torch.float64, requires_grad=True)
    ref = O.rnn_direction(x64, lens, cp, reverse, cell="relu_plain", sbn="none")
    (ref * g.double().cpu()).sum().backward()
    assert _rel(out.detach().cpu().numpy(), ref.detach().numpy()) < 1e-4
    assert _rel(x.grad.cpu().numpy(), x64.grad.numpy()) < 1e-3
    gb = blob.grad.cpu().numpy()
    assert _rel(gb[:H * I].reshape(H, I), cp["W"].grad.numpy()) < 1e-3
    assert _rel(gb[H * I:H * I + H * H].reshape(H, H), cp["U"].grad.numpy()) < 1e-3
    assert _rel(gb[H * I + H * H:H * I + H * H + H], cp["B"].grad.numpy()) < 1e-3      # b_W and b_U see the same gradient
    assert _rel(gb[H * I + H * H + H:], cp["B"].grad.numpy()) < 1e-3
    # the per-step call of the same blob (what the reference does, one call per timestep) gives the same states
    if not reverse:
        h = torch.zeros((B, H), device="cuda")
        with torch.no_grad():
            for t in range(3):                         # all utterances but one are still running at t < 1; compare row 0
                o, h = cell.model(input_data=x[t], input_h=h, params=blob)
                assert _rel(o[0].cpu().numpy(), out[t, 0].detach().cpu().numpy()) < 1e-5


def test_engine_mkldnn_rnn_model_runs_through_the_blob():
    from deepspeech_b200 import deepSpeech_NCHW as ds
    from deepspeech_b200.deepSpeech_dummy import DummyData
    params = SimpleNamespace(use_fp16=False, num_filters=4, batch_size=4, num_hidden=32, num_rnn_layers=2, engine="mkldnn_rnn",
                             rnn_type="bi-dir", wiring="as_written", seed=3)
    m = ds.model(params)
    assert len(m.rnn_blob) == 4 and len(m.rnn_W) == 0
    feats, idx, vals, shape, seq_lens = DummyData(4, seed=1).inputs(bucket=0)
    m.zero_grad(set_to_none=True)
    logits = ds.inference(None, torch.from_numpy(feats).cuda(), seq_lens, params)
    loss = ds.loss(logits, (idx, vals, shape), seq_lens)
    loss.backward()
    assert np.isfinite(loss.item())
    assert m.rnn_blob[2].grad is not None and torch.isfinite(m.rnn_blob[2].grad).all()     # the live (last) layer
    assert m.rnn_blob[0].grad is None                                                     # dead layers get no gradient


def test_linear_and_custom_rnn_cell_positive_paths():
    from deepspeech_b200 import custom_ops as C
    rng = np.random.default_rng(5)
    B, I, H = 6, 10, 14
    x = torch.tensor(rng.standard_normal((B, I)), dtype=torch.float32, device="cuda")
    h = torch.tensor(rng.standard_normal((B, H)), dtype=torch.float32, device="cuda")
    y = C._linear([x, h], H, True, scope="t_lin_pos")
    M = C._get_variable("t_lin_pos", "Matrix", (H, I + H), None, True, "cuda").detach().cpu().numpy()
    bias = C._get_variable("t_lin_pos", "Bias", (H,), 0.0, True, "cuda").detach().cpu().numpy()
    ref = np.concatenate([x.cpu().numpy(), h.cpu().numpy()], 1).astype(np.float64) @ M.T.astype(np.float64) + bias
    assert y.shape == (B, H) and _rel(y.detach().cpu().numpy(), ref) < 1e-5
    y1 = C._linear(x, H, False, scope="t_lin_pos1")                      # a single tensor instead of a list, no bias
    M1 = C._get_variable("t_lin_pos1", "Matrix", (H, I), None, True, "cuda").detach().cpu().numpy()
    assert _rel(y1.detach().cpu().numpy(), x.cpu().numpy().astype(np.float64) @ M1.T) < 1e-5
    cell = C.CustomRNNCell(H)                                            # relu6(_linear([x, h]))  (src/custom_ops.py:27-33)
    out, state = cell(x, h, scope="t_cell_pos")
    Mc = C._get_variable("t_cell_pos", "Matrix", (H, I + H), None, True, "cuda").detach().cpu().numpy()
    refc = np.clip(np.concatenate([x.cpu().numpy(), h.cpu().numpy()], 1).astype(np.float64) @ Mc.T, 0.0, 6.0)
    assert out is state and _rel(out.detach().cpu().numpy(), refc) < 1e-5
    assert cell.output_size == H and cell.zero_state(3).shape == (3, H)


def test_device_dummy_producer_matches_host_generator_bit_for_bit():
    """the device-side dummy batch producer against the host generator with the same seed - features,
    SparseTensor indices / values, shapes and lengths identical over a walk through the bucket schedule."""
    from deepspeech_b200.deepSpeech_dummy import DeviceDummyData, DummyData
    host, dev = DummyData(4, seed=11), DeviceDummyData(4, seed=11)
    for step in range(40):
        f0, i0, v0, s0, l0 = host.inputs()
        f1, i1, v1, s1, l1 = dev.inputs()
        assert f1.is_cuda and i1.is_cuda and v1.is_cuda
        assert np.array_equal(f0, f1.cpu().numpy()) and np.array_equal(i0, i1.cpu().numpy()) and np.array_equal(v0, v1.cpu().numpy())
        assert np.array_equal(s0, s1) and np.array_equal(l0, l1)
    assert host.current == dev.current


def test_ctc_loss_with_device_labels_matches_host_path_and_flags_bad_rows():
    from deepspeech_b200 import ctc
    rng = np.random.default_rng(8)
    T, B, K, L = 40, 6, 29, 9
    x = torch.tensor(rng.standard_normal((T, B, K)), dtype=torch.float32, device="cuda")
    dense = rng.integers(0, 28, size=(B, L))
    dense[:, -1] = 28                                                   # the dummy generator's terminator
    keep = dense != 0
    idx = np.argwhere(keep).astype(np.int64)
    vals = dense[keep].astype(np.int32)
    sl = np.full(B, T, np.int32)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    la = ctc.ctc_loss((idx, vals, np.asarray([B, L])), xa, sl, preprocess_collapse_repeated=True)
    lb = ctc.ctc_loss((torch.from_numpy(idx).cuda(), torch.from_numpy(vals).cuda(), np.asarray([B, L])), xb, sl, preprocess_collapse_repeated=True)
    la.sum().backward(); lb.sum().backward()
    assert torch.equal(la, lb) and torch.equal(xa.grad, xb.grad)
    # a row with too few frames: the host path raises as TF does, the device path answers inf and a zero gradient row
    sl2 = sl.copy(); sl2[2] = 3
    with pytest.raises(ValueError):
        ctc.ctc_loss((idx, vals, np.asarray([B, L])), x, sl2, preprocess_collapse_repeated=True)
    xc = x.clone().requires_grad_(True)
    lc = ctc.ctc_loss((torch.from_numpy(idx).cuda(), torch.from_numpy(vals).cuda(), np.asarray([B, L])), xc, sl2, preprocess_collapse_repeated=True)
    ok = torch.ones(B, dtype=torch.bool, device="cuda"); ok[2] = False
    lc[ok].sum().backward()
    assert torch.isinf(lc[2]) and torch.isfinite(lc[ok]).all() and torch.equal(lc[ok], la[ok].detach())
    assert xc.grad[:, 2].abs().max().item() == 0.0
