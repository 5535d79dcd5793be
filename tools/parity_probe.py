#!/usr/bin/env python
"""Where does the bf16 path's deviation from the fp64 oracle come from?  (needs a GPU)
Runs the train.sh model (T=1500 bucket) layer by layer under a few variants and prints the max-norm error after every
layer and at the logits:  usage: tools/parity_probe.py [variant ...]
  base        the shipped bf16 path
  fe32        fp32 front-end (direct conv kernels), bf16 recurrent stack
  w16         oracle run on bf16-ROUNDED weights (isolates the weight rounding)
  slab32      DS2_RNN_SLAB_F32=1 (fp32 partial sums across the cluster)
  sum16       layer output fw + bw formed from the bf16-rounded hidden states (round 1 behaviour)"""
import os
import sys
import time
from types import SimpleNamespace

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepspeech_b200 import deepSpeech_NCHW as ds, ops
from deepspeech_b200._lib import lib
from oracle import ds2_oracle as O

variants = sys.argv[1:] or ["base"]
torch.set_num_threads(os.cpu_count() or 1)
B, H, L = 32, 1760, 7
rng = np.random.default_rng(1234)
p, layers = O.init_params(rng, num_filters=32, num_hidden=H, num_layers=L)
feats, idx, vals, shape, seq_lens = O.dummy_batch(rng, B, 14)


def err(got, ref):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    return np.abs(got - ref).max() / np.abs(ref).max(), np.sqrt(((got - ref) ** 2).mean()) / np.abs(ref).max()


def oracle(round_w):
    bf = (lambda t: t.float().to(torch.bfloat16).double()) if round_w else (lambda t: t)
    with torch.no_grad():
        x = O.conv_frontend(torch.tensor(feats, dtype=torch.float64), p, "NCHW")
        lens = O.get_rnn_seqlen(seq_lens)
        outs = [x.numpy()]
        for lp in layers:
            q = {d: {"W": bf(lp[d]["W"]), "U": bf(lp[d]["U"]), "B": lp[d]["B"]} for d in ("fw", "bw")}
            x = O.rnn_direction(x, lens, q["fw"], False) + O.rnn_direction(x, lens, q["bw"], True)
            outs.append(x.numpy())
        outs.append(O.fc_logits(x, bf(p["softmax_linear.weights"].double()), p["softmax_linear.biases"].double()).numpy())
    return outs, lens


t0 = time.time()
ref_full, lens = oracle(False)
ref_w16 = oracle(True)[0] if "w16" in variants else None
print("oracle: %.0f s" % (time.time() - t0))
lens_d = torch.from_numpy(np.asarray(lens, np.int32)).cuda()
for v in variants:
    ops.USE_FP32_LAYER_SUM = v not in ("sum16",)
    if v == "slab32":
        os.environ["DS2_RNN_SLAB_F32"] = "1"
    else:
        os.environ.pop("DS2_RNN_SLAB_F32", None)
    lib.ds2_debug_reload_env()
    params = SimpleNamespace(use_fp16=True, num_filters=32, batch_size=B, num_hidden=H, num_rnn_layers=L, engine="mkl",
                             rnn_type="bi-dir", wiring="stacked", seed=0)
    m = ds.model(params)
    m.load_oracle_params(p, layers)
    ref = ref_w16 if v == "w16" else ref_full
    with torch.no_grad():
        f = torch.from_numpy(feats).cuda()
        if v == "fe32":
            m.use_bf16 = False
            x = m.frontend(f).to(torch.bfloat16)
            m.use_bf16 = True
        else:
            x = m.frontend(f)
        rep = ["fe %.1e/%.1e" % err(x.float().cpu().numpy(), ref[0])]
        for l in range(L):
            x = m._layer_eval(x, l, lens_d)
            rep.append("L%d %.1e/%.1e" % ((l,) + err(x.float().cpu().numpy(), ref[l + 1])))
        T2 = x.shape[0]
        logits = ops.linear(x.reshape(T2 * B, H), m.fc_w, m.fc_b, True).view(T2, B, 29)
        rep.append("logits %.2e/%.1e" % err(logits.float().cpu().numpy(), ref[L + 1]))
    print("%-7s (max/rms of scale): %s" % (v, ", ".join(rep)))
