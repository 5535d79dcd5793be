"""Optimizer step: schedule / EMA decay on the CPU, the fused kernel against the fp64 oracle on the GPU."""
import numpy as np
import pytest
import torch

from oracle import ds2_oracle as O


def test_learning_rate_schedule_and_ema_decay():
    # staircase: constant inside a decay period, one factor per period (src/deepSpeech_train.py:231-250)
    assert O.decayed_learning_rate(1e-5, 0, 100, 0.9) == 1e-5
    assert O.decayed_learning_rate(1e-5, 99, 100, 0.9) == 1e-5
    assert abs(O.decayed_learning_rate(1e-5, 250, 100, 0.9) - 1e-5 * 0.81) < 1e-20
    # EMA warm-up of tf.train.ExponentialMovingAverage(decay, num_updates)
    assert abs(O.ema_decay(0.9999, 1) - 2.0 / 11.0) < 1e-15
    assert O.ema_decay(0.9999, 10 ** 9) == 0.9999


def test_adam_oracle_against_torch_adam():
    """The restated Adam equals torch.optim.Adam (same published algorithm) up to where epsilon enters:
    TF adds eps to sqrt(v) BEFORE bias correction ("epsilon hat"), torch after; with eps -> 0 they coincide."""
    rng = np.random.default_rng(0)
    w0 = rng.standard_normal(50)
    p = torch.tensor(w0, dtype=torch.float64, requires_grad=True)
    opt = torch.optim.Adam([p], lr=1e-3, eps=1e-30)
    w, m, v = w0.copy(), np.zeros(50), np.zeros(50)
    for step in range(1, 6):
        g = rng.standard_normal(50)
        p.grad = torch.tensor(g)
        opt.step()
        w, m, v, _ = O.adam_ema_step_np(w, g, m, v, None, step, 1e-3, 0.9999, eps=1e-30)
        assert np.abs(w - p.detach().numpy()).max() < 1e-12


@pytest.mark.gpu
def test_fused_adam_ema_kernel_vs_oracle():
    from deepspeech_b200.optim import AdamEMA, set_learning_rate
    rng = np.random.default_rng(3)
    shapes = [(7,), (4096,), (33, 129), (1, 1), (5000,)]          # odd sizes, a tensor of exactly one chunk, several chunks
    ws = [rng.standard_normal(s) for s in shapes]
    params = [torch.tensor(w, dtype=torch.float32, device="cuda", requires_grad=True) for w in ws]
    lr_fn, decay_steps = set_learning_rate(1e-2, 32, 0.001, 0.5, num_per_epoch=64000)    # decay every 2 steps
    assert decay_steps == 2
    opt = AdamEMA(params, lr_fn, moving_avg_decay=0.9)
    ref = [dict(w=np.asarray(p.detach().cpu(), np.float64), m=np.zeros(s), v=np.zeros(s), sh=np.asarray(p.detach().cpu(), np.float64))
           for p, s in zip(params, shapes)]
    for step in range(1, 6):
        lr = O.decayed_learning_rate(1e-2, step - 1, 2, 0.5)
        assert abs(opt.learning_rate() - lr) < 1e-12
        for p, r in zip(params, ref):
            g = rng.standard_normal(tuple(p.shape)).astype(np.float32)
            p.grad = torch.tensor(g, device="cuda")               # a fresh tensor every step: the pointer table is rebuilt
            r["w"], r["m"], r["v"], r["sh"] = O.adam_ema_step_np(r["w"], g, r["m"], r["v"], r["sh"], step, lr, 0.9, grad_scale=0.5)
        opt.step(grad_scale=0.5)
        torch.cuda.synchronize()
        for i, (p, r) in enumerate(zip(params, ref)):
            assert np.abs(p.detach().cpu().numpy() - r["w"]).max() < 2e-6, (step, i)
            assert np.abs(opt.shadow[i].cpu().numpy() - r["sh"]).max() < 2e-6, (step, i)
            assert np.abs(opt.m[i].cpu().numpy() - r["m"]).max() < 1e-6
            assert np.abs(opt.v[i].cpu().numpy() - r["v"]).max() < 1e-6
    with pytest.raises(RuntimeError):
        AdamEMA([torch.zeros(3, requires_grad=True)], 1e-3)      # CPU tensor: no fallback
