"""BASELINE-shape parity against the fp64 ORACLE (not against this repository's own fp32 kernels).

north_star's bars, each written where it is asserted:
  * bf16 recurrence vs the fp64 reference restatement: <= 1e-2 of the output / logits scale;
  * CTC loss <= 1e-4 relative, CTC gradient <= 1e-3 of its scale, fp32.
The oracle (oracle/ds2_oracle.py, torch-CPU fp64) needs tens of seconds at these sizes; every test prints the time it
took.  Shapes: train.sh layer (T2=366, B=32, H=1760, ragged lengths), the whole 7-layer train.sh model on the
T=1500 dummy bucket (BASELINE configs 1 and 4(i)), the CTC sweep (config 2), the config-3 GRU layer (H=800)."""
import os
import time
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import ds2_oracle as O

pytestmark = pytest.mark.gpu


def _threads():
    torch.set_num_threads(max(1, os.cpu_count() or 1))


def _bf(a):
    return torch.tensor(np.asarray(a), dtype=torch.float32).to(torch.bfloat16).double()


def _scale_err(got, ref):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))


def test_relu_layer_train_sh_shape_vs_oracle():
    """One bi-directional clipped-ReLU layer at the train.sh shape through ds2_rnn_relu_fwd / _bwd (persistent bf16
    kernels) against O.rnn_direction on the same bf16-rounded weights and inputs.
    Forward: <= 1e-2 of the output scale (north_star).  Backward: d loss / d pre-activation compared ELEMENT-WISE on the
    entries whose oracle pre-activation is farther from both kinks (0 and 20) than twice the largest forward deviation,
    i.e. whose ReLU mask cannot differ between the two computations.  A flipped mask elsewhere still perturbs those
    entries through U^T, so the bar is statistical: median <= 3e-3 and 99th percentile <= 5e-2 of the gradient scale."""
    from deepspeech_b200 import ops
    from deepspeech_b200._lib import DS2_BF16, check, lib
    _threads()
    T, B, H = 366, 32, 1760
    I = H
    rng = np.random.default_rng(2024)
    lp = {}
    for d in ("fw", "bw"):
        lp[d] = {"W": _bf(rng.standard_normal((H, I)) * (1.0 / np.sqrt(I))).requires_grad_(True),
                 "U": _bf(rng.standard_normal((H, H)) * (0.6 / np.sqrt(H))).requires_grad_(True),
                 "B": torch.tensor(rng.standard_normal(H) * 0.1).float().double().requires_grad_(True)}
    x64 = _bf(rng.standard_normal((T, B, I)))
    lens = [T] + [int(v) for v in rng.integers(T // 3, T + 1, size=B - 1)]
    g64 = torch.tensor(rng.standard_normal((T, B, H))).float().double()
    t0 = time.time()
    zf, zb = [None] * T, [None] * T
    ref = O.rnn_direction(x64, lens, lp["fw"], False, preacts=zf) + O.rnn_direction(x64, lens, lp["bw"], True, preacts=zb)
    (ref * g64).sum().backward()
    t_oracle = time.time() - t0
    ref_np = ref.detach().numpy()

    dev = "cuda"
    p, st = ops._p, ops._stream
    xg = x64.to(torch.bfloat16).to(dev)
    Wg = torch.cat([lp["fw"]["W"], lp["bw"]["W"]], 0).detach().to(torch.bfloat16).to(dev)
    Ug = torch.stack([lp["fw"]["U"], lp["bw"]["U"]], 0).detach().to(torch.bfloat16).to(dev).contiguous()
    bias = torch.stack([lp["fw"]["B"], lp["bw"]["B"]], 0).detach().float().to(dev).contiguous()
    lens_d = torch.tensor(lens, dtype=torch.int32, device=dev)
    wx = ops.gemm_nt(xg.view(T * B, I), Wg)
    hs = torch.empty((2, T, B, H), dtype=torch.bfloat16, device=dev)
    wsb = lib.ds2_rnn_workspace_bytes(T, B, H, 2, DS2_BF16)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    check(lib.ds2_rnn_relu_fwd(p(wx), ops.SBN_SCALE, p(Ug), p(bias), p(lens_d), p(hs), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "fwd")
    out = (hs[0].float() + hs[1].float()).cpu().numpy()
    err_f = _scale_err(out, ref_np)
    for b, n in enumerate(lens):
        assert np.abs(out[n:, b]).max(initial=0.0) == 0.0
    # backward on the kernel's own stored activations
    dz = torch.empty((T, B, 2 * H), dtype=torch.bfloat16, device=dev)
    db = torch.empty((2, H), dtype=torch.float32, device=dev)
    gdev = g64.float().to(dev).contiguous()
    check(lib.ds2_rnn_relu_bwd(p(gdev), p(Ug), p(lens_d), p(hs), p(dz), p(db), DS2_BF16, T, B, H, 2, 2, 20.0, p(ws), wsb, st()), "bwd")
    torch.cuda.synchronize()
    dz_np = dz.float().cpu().numpy().reshape(T, B, 2, H)
    stats = []
    fwd_dev = np.abs(out - ref_np).max()
    for d, zl in enumerate((zf, zb)):
        z = torch.stack([v.detach() for v in zl], 0).numpy()
        dz_ref = torch.stack([v.grad if v.grad is not None else torch.zeros_like(v) for v in zl], 0).numpy()
        valid = np.zeros((T, B, 1), bool)
        for b, n in enumerate(lens):
            valid[:n, b] = True
        margin = 2.0 * fwd_dev
        safe = valid & (np.abs(z) > margin) & (np.abs(z - 20.0) > margin)
        e = np.abs(dz_np[:, :, d] - dz_ref) / np.abs(dz_ref).max()
        es = e[np.broadcast_to(safe, e.shape)]
        stats.append((float(np.median(es)), float(np.percentile(es, 99)), float(es.max()), float(safe.mean())))
        cos = float((dz_np[:, :, d] * dz_ref).sum() / (np.linalg.norm(dz_np[:, :, d]) * np.linalg.norm(dz_ref)))
        assert cos > 0.99, cos
    print("relu layer T=%d B=%d H=%d: fwd err %.2e of scale; dz safe-entry err median/p99/max (fw) %.1e/%.1e/%.1e, (bw) %.1e/%.1e/%.1e, "
          "safe fraction %.2f; oracle %.1f s" % (T, B, H, err_f, *stats[0][:3], *stats[1][:3], stats[0][3], t_oracle))
    assert err_f < 1e-2, err_f                                   # north_star: bf16 recurrence <= 1e-2
    for med, p99, _, _ in stats:
        assert med < 3e-3 and p99 < 5e-2, (med, p99)


def _oracle_stack(feats, seq_lens, p, layers, round_weights):
    """O.model_forward layer by layer (so that the error after every layer can be reported); with round_weights the
    recurrent and FC weights are rounded to bf16 first - the operand precision the tensor-core path uses by design."""
    bf = (lambda t: t.float().to(torch.bfloat16).double()) if round_weights else (lambda t: t)
    with torch.no_grad():
        x = O.conv_frontend(torch.tensor(feats, dtype=torch.float64), p, "NCHW")
        lens = O.get_rnn_seqlen(seq_lens)
        outs = [x.numpy()]
        for lp in layers:
            q = {d: {"W": bf(lp[d]["W"]), "U": bf(lp[d]["U"]), "B": lp[d]["B"]} for d in ("fw", "bw")}
            x = O.rnn_direction(x, lens, q["fw"], False) + O.rnn_direction(x, lens, q["bw"], True)
            outs.append(x.numpy())
        logits = O.fc_logits(x, bf(p["softmax_linear.weights"].double()), p["softmax_linear.biases"].double())
    return outs, logits, lens


def test_train_sh_model_logits_and_loss_vs_oracle():
    """BASELINE configs 1 / 4(i): the 7-layer bi-RNN H=1760 train.sh model on the T=1500 dummy bucket, bf16 tensor-core
    path through the reference-facing API, against the fp64 oracle, layer by layer.
      * oracle on the bf16-ROUNDED recurrent/FC weights (what the kernels are given): logits <= 1e-2 of their scale
        (north_star's bar for the bf16 recurrence), loss <= 1e-2 relative;
      * oracle on the fp32 MASTER weights: measured 1.1e-2 of the logits' scale (rms 2.3e-3) - a fifth of it is the weight
        rounding itself, the rest the bf16 state fed back through 7 x 366 steps of a model that random initialisation puts
        deep into saturation (|logit| up to 500).  Asserted at 1.5e-2 and reported; DESIGN.md section 2 has the break-down
        (tools/parity_probe.py) and what was tried (fp32 layer sums: kept; fp32 partial-sum exchange: -10 % error, +17 % time)."""
    from deepspeech_b200 import deepSpeech_NCHW as ds
    _threads()
    B, H, L = 32, 1760, 7
    rng = np.random.default_rng(1234)
    p, layers = O.init_params(rng, num_filters=32, num_hidden=H, num_layers=L)
    feats, idx, vals, shape, seq_lens = O.dummy_batch(rng, B, 14)                       # T = 1500 frames
    params = SimpleNamespace(use_fp16=True, num_filters=32, batch_size=B, num_hidden=H, num_rnn_layers=L, engine="mkl",
                             rnn_type="bi-dir", wiring="stacked", seed=0)
    m = ds.model(params)
    m.load_oracle_params(p, layers)
    with torch.no_grad():
        logits = ds.inference(None, torch.from_numpy(feats).cuda(), seq_lens, params)
        loss = ds.loss(logits, (idx, vals, shape), seq_lens)
        x_gpu = m.frontend(torch.from_numpy(feats).cuda())      # the same stack layer by layer, for the per-layer report
    torch.cuda.synchronize()
    t0 = time.time()
    outs16, logits16, lens = _oracle_stack(feats, seq_lens, p, layers, True)
    outs32, logits32, _ = _oracle_stack(feats, seq_lens, p, layers, False)
    labels = O.sparse_to_label_seqs(idx, vals, shape)
    with torch.no_grad():
        loss16 = O.ctc_loss_torch(logits16, labels, lens).mean().item()
        loss32 = O.ctc_loss_torch(logits32, labels, lens).mean().item()
    t_oracle = time.time() - t0
    report = ["frontend %.1e" % _scale_err(x_gpu.float().cpu().numpy(), outs16[0])]
    cur_gpu = x_gpu
    lens_d = torch.from_numpy(np.asarray(lens, np.int32)).cuda()
    for l in range(L):
        cur_gpu = m._layer_eval(cur_gpu, l, lens_d)
        g = cur_gpu.float().cpu().numpy()
        report.append("layer %d %.1e (%.1e)" % (l, _scale_err(g, outs16[l + 1]), _scale_err(g, outs32[l + 1])))
    lg = logits.float().cpu().numpy()
    err16, err32 = _scale_err(lg, logits16.numpy()), _scale_err(lg, logits32.numpy())
    el16, el32 = abs(loss.item() - loss16) / abs(loss16), abs(loss.item() - loss32) / abs(loss32)
    print("train.sh model T=1500: logits err %.2e of scale %.0f vs the oracle on bf16-rounded weights, %.2e on fp32 weights; loss %.2f "
          "vs %.2f / %.2f (rel %.1e / %.1e); per layer, rounded (fp32) weights: %s; oracle %.0f s"
          % (err16, float(logits16.abs().max()), err32, loss.item(), loss16, loss32, el16, el32, ", ".join(report), t_oracle))
    assert err16 < 1e-2, err16                                   # north_star: bf16 recurrence <= 1e-2 on logits
    assert el16 < 1e-2 and el32 < 1e-2, (el16, el32)
    assert err32 < 1.5e-2, err32                                 # fp32 master weights: see the docstring


@pytest.mark.parametrize("T,L,B", [(200, 50, 64), (800, 200, 64), (1600, 400, 64)])
def test_ctc_sweep_vs_oracle(T, L, B):
    """BASELINE config 2 against torch's fp64 CTC (the independent implementation the oracle is pinned to) and the
    oracle's own numpy restatement on two rows: loss <= 1e-4 relative, gradient <= 1e-3 of its scale; ragged."""
    from deepspeech_b200 import ops
    _threads()
    rng = np.random.default_rng(7)
    x = rng.standard_normal((T, B, 29)).astype(np.float32)
    labs = []
    for b in range(B):
        r = rng.integers(0, 28, size=int(rng.integers(L // 2, L + 1)))
        labs.append([int(v) for i, v in enumerate(r) if i == 0 or v != r[i - 1]])
    Lm = max(len(r) for r in labs)
    lab = np.zeros((B, Lm), np.int32)
    for b, r in enumerate(labs):
        lab[b, :len(r)] = r
    ll = np.asarray([len(r) for r in labs], np.int32)
    sl = rng.integers(T // 2 + Lm, T + 1, size=B).astype(np.int32)
    sl[0] = T
    xt = torch.tensor(x, device="cuda", requires_grad=True)
    loss = ops.ctc_loss_dense(xt, lab, ll, np.zeros(B, np.int32), sl)
    loss.sum().backward()
    t0 = time.time()
    x64 = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    ref = torch.nn.functional.ctc_loss(torch.log_softmax(x64, 2), torch.tensor(sum(labs, []), dtype=torch.long), torch.tensor(sl, dtype=torch.long),
                                       torch.tensor(ll, dtype=torch.long), blank=28, reduction="none")
    ref.sum().backward()
    l2, g2 = O.ctc_loss_and_grad_np(x[:, :2].astype(np.float64), labs[:2], sl[:2])
    t_oracle = time.time() - t0
    el = float(((loss.detach().cpu().double() - ref.detach()).abs() / ref.detach().abs()).max())
    eg = _scale_err(xt.grad.cpu().numpy(), x64.grad.numpy())
    print("CTC (T=%d, L<=%d, B=%d): loss rel err %.2e, grad err %.2e of scale; oracle %.1f s" % (T, L, B, el, eg, t_oracle))
    assert el < 1e-4, el
    assert eg < 1e-3, eg
    assert np.abs(loss.detach().cpu().numpy()[:2] - l2).max() / np.abs(l2).max() < 1e-4
    assert _scale_err(xt.grad.cpu().numpy()[:, :2], g2) < 1e-3


def test_gru_layer_config3_shape_vs_oracle():
    """BASELINE config 3 shape (bi-GRU H=800, B=32, T2=366, ragged): persistent bf16 kernels against
    O.rnn_direction(cell='gru') on the same bf16-rounded weights; <= 1e-2 of the output scale (|h| <= 1)."""
    from deepspeech_b200 import ops
    _threads()
    T, B, H = 366, 32, 800
    I = H
    rng = np.random.default_rng(33)
    lp = {}
    for d in ("fw", "bw"):
        lp[d] = {"W": _bf(rng.standard_normal((3 * H, I)) * (1.0 / np.sqrt(I))), "U": _bf(rng.standard_normal((3 * H, H)) * (0.8 / np.sqrt(H))),
                 "BW": torch.tensor(rng.standard_normal(3 * H) * 0.1).float().double(),
                 "BU": torch.tensor(rng.standard_normal(3 * H) * 0.1).float().double()}
    x64 = _bf(rng.standard_normal((T, B, I)))
    lens = [T] + [int(v) for v in rng.integers(T // 3, T + 1, size=B - 1)]
    t0 = time.time()
    with torch.no_grad():
        ref = (O.rnn_direction(x64, lens, lp["fw"], False, cell="gru") + O.rnn_direction(x64, lens, lp["bw"], True, cell="gru")).numpy()
    t_oracle = time.time() - t0
    dev = "cuda"
    W = torch.cat([lp["fw"]["W"], lp["bw"]["W"]], 0).float().to(dev)
    U = torch.stack([lp["fw"]["U"], lp["bw"]["U"]], 0).float().to(dev)
    bw = torch.stack([lp["fw"]["BW"], lp["bw"]["BW"]], 0).float().to(dev)
    bu = torch.stack([lp["fw"]["BU"], lp["bw"]["BU"]], 0).float().to(dev)
    with torch.no_grad():
        out = ops.bigru_layer(x64.to(torch.bfloat16).to(dev), W, U, bw, bu, torch.tensor(lens, dtype=torch.int32, device=dev), 0b10, True,
                              torch.float32)
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    err = _scale_err(o, ref)
    print("GRU layer T=%d B=%d H=%d: err %.2e of scale %.2f; oracle %.1f s" % (T, B, H, err, np.abs(ref).max(), t_oracle))
    assert err < 1e-2, err
    for b, n in enumerate(lens):
        assert np.abs(o[n:, b]).max(initial=0.0) == 0.0
