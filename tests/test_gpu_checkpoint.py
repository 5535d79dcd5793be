"""Checkpoint round trip on the device: train, save (weights + Adam slots + EMA shadows + step +
deepSpeech_parameters.json), restore into a freshly built model / optimizer, and the NEXT training step is bit-identical
to the one the original run takes (src/deepSpeech_train.py:354-356, :383-398, :511-513; EMA restore for evaluation:
src/deepSpeech_test.py:218-220)."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_save_restore_then_next_step_is_bit_identical(tmp_path):
    from deepspeech_b200 import checkpoint as C, deepSpeech_NCHW as ds
    from deepspeech_b200.deepSpeech_dummy import DummyData
    from deepspeech_b200.model import DeepSpeech2
    from deepspeech_b200.optim import AdamEMA
    mk = lambda seed: DeepSpeech2(num_filters=4, num_hidden=32, num_rnn_layers=2, layout="NCHW", wiring="stacked", use_bf16=False, seed=seed)
    data = DummyData(4, seed=3)
    batches = [data.inputs(bucket=0) for _ in range(3)]
    params = SimpleNamespace(num_hidden=32, num_rnn_layers=2, rnn_type="bi-dir", num_filters=4, initial_lr=1e-3, temporal_stride=2)

    def step(m, opt, batch):
        feats, idx, vals, shape, seq_lens = batch
        m.zero_grad(set_to_none=True)
        logits = m(torch.from_numpy(feats).cuda(), seq_lens)
        loss = ds.loss(logits, (idx, vals, shape), seq_lens)
        loss.backward()
        opt.step()
        return loss.item()

    m = mk(1)
    opt = AdamEMA(m.parameters(), 1e-3)
    step(m, opt, batches[0]); step(m, opt, batches[1])
    d = str(tmp_path)
    C.save_parameters_json(d, params)
    C.save(d, m, opt)
    saved = [p.detach().clone() for p in m.parameters()], [t.clone() for t in opt.m], [t.clone() for t in opt.v], [t.clone() for t in opt.shadow]
    want = step(m, opt, batches[2])
    m2 = mk(99)                                        # different initial weights: everything must come from the file
    opt2 = AdamEMA(m2.parameters(), 1e-3)
    assert C.restore(d, m2, opt2) == 2 and C.load_parameters_json(d).num_hidden == 32
    # the restored state is the saved state, bit for bit
    for a, b in zip(saved[0], m2.parameters()):
        assert torch.equal(a, b)
    for i in range(len(opt2.params)):
        assert torch.equal(saved[1][i], opt2.m[i]) and torch.equal(saved[2][i], opt2.v[i]) and torch.equal(saved[3][i], opt2.shadow[i])
    # ... so the next step computes the same loss exactly; its gradients are sums accumulated with atomics (weight-gradient
    # and batch-norm reductions), whose order differs from run to run: the updated weights agree to fp32 round-off
    got = step(m2, opt2, batches[2])
    assert got == want
    for (name, a), b in zip(m.named_parameters(), m2.parameters()):
        if name in ("conv1_b", "conv2_b"):
            # a bias in front of a batch norm has an analytically ZERO gradient: what Adam divides by sqrt(v) there is pure
            # round-off, so its update is +-lr with a sign that the summation order decides
            assert torch.allclose(a, b, rtol=0.0, atol=2.5e-3), name
        else:
            assert torch.allclose(a, b, rtol=1e-3, atol=2e-5), name
    m3 = mk(5)
    C.restore(d, m3, use_moving_averages=True)
    z = np.load(C.latest_checkpoint(d))
    for name, p in m3.named_parameters():
        assert np.array_equal(p.detach().cpu().numpy(), z[name + C.EMA_SUFFIX])
