#!/usr/bin/env python
"""Error of the bf16 recurrent layer against the fp64 oracle on bf16-rounded weights (needs a GPU): max abs error over the
output scale, for a few shapes.  Used to judge numerics-affecting kernel options (set them through the environment)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepspeech_b200 import ops
from oracle import ds2_oracle as O

for (T, B, I, H) in [(24, 32, 64, 800), (12, 32, 256, 1760), (60, 32, 128, 512)]:
    rng = np.random.default_rng(T * 13 + H)
    bf = lambda a: torch.tensor(a, dtype=torch.float32).to(torch.bfloat16).double()
    lp = {}
    for d in ("fw", "bw"):
        lp[d] = {"W": bf(rng.standard_normal((H, I)) * (1.0 / np.sqrt(I))), "U": bf(rng.standard_normal((H, H)) * (0.7 / np.sqrt(H))),
                 "B": torch.tensor(rng.standard_normal(H) * 0.1).float().double()}
    x64 = bf(rng.standard_normal((T, B, I)))
    lens = [T] * B
    ref = O.rnn_direction(x64, lens, lp["fw"], False) + O.rnn_direction(x64, lens, lp["bw"], True)
    dev = "cuda"
    x = x64.to(torch.bfloat16).to(dev)
    W = torch.cat([lp["fw"]["W"], lp["bw"]["W"]], 0).float().to(dev)
    U = torch.stack([lp["fw"]["U"], lp["bw"]["U"]], 0).float().to(dev)
    Bv = torch.stack([lp["fw"]["B"], lp["bw"]["B"]], 0).float().to(dev)
    out = ops.birnn_layer(x, W, U, Bv, torch.tensor(lens, dtype=torch.int32, device=dev), 0b10, True, torch.float32)
    o = out.cpu().double().numpy()
    r = ref.numpy()
    err = np.abs(o - r)
    print("T=%d B=%d I=%d H=%d: max|err|/max|ref| = %.3e   rms err / rms ref = %.3e   (last step alone: %.3e)"
          % (T, B, I, H, err.max() / np.abs(r).max(), np.sqrt((err ** 2).mean()) / np.sqrt((r ** 2).mean()),
             np.abs(o[-1] - r[-1]).max() / np.abs(r).max()))
