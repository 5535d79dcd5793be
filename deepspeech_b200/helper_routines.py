"""Initialisers of the reference's variable helpers (src/helper_routines.py:31-78), host side, seeded.

`_variable_on_cpu` pinned TF variables to /cpu; here parameters live on the current CUDA device and
these helpers only reproduce the *initial values* so that "identical weights" is reproducible."""
from __future__ import annotations

import math

import numpy as np
import torch


def variance_scaling(rng: np.random.Generator, shape, factor: float = 2.0) -> np.ndarray:
    """tf.contrib.layers.variance_scaling_initializer(factor=2.0, mode='FAN_IN', uniform=False)
    (src/helper_routines.py:66-70): truncated normal (|z| <= 2) with stddev sqrt(1.3*factor/fan_in);
    fan_in = prod(shape[:-1]) for conv filters, shape[-2] for matrices."""
    fan_in = float(np.prod(shape[:-1])) if len(shape) == 4 else float(shape[-2])
    std = math.sqrt(1.3 * factor / fan_in)
    out = rng.standard_normal(shape)
    bad = np.abs(out) > 2.0
    while bad.any():
        out[bad] = rng.standard_normal(int(bad.sum()))
        bad = np.abs(out) > 2.0
    return (out * std).astype(np.float32)


def stable_seed(name: str, base: int = 1234) -> int:
    """Seed derived from a variable's scope name, identical in every process and on every data-parallel rank
    (Python randomises str hashes per process; crc32 does not)."""
    import zlib
    return (zlib.crc32(name.encode("utf-8")) + int(base)) % (2 ** 32)


def glorot_uniform(rng: np.random.Generator, shape) -> np.ndarray:
    """tf.get_variable's default initializer for the RNN W/U (src/custom_ops.py:60,66)."""
    lim = math.sqrt(6.0 / (shape[-2] + shape[-1]))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def _variable_on_cpu(name, shape, initializer=None, use_fp16=False, trainable=True, device="cuda"):
    """Same argument list as src/helper_routines.py:31; returns a torch Parameter on `device`.
    `initializer` is a callable shape -> ndarray, a float constant, or None (glorot uniform)."""
    if isinstance(shape, int):
        shape = [shape]
    if initializer is None:
        val = glorot_uniform(np.random.default_rng(stable_seed(name)), tuple(shape)) if len(shape) >= 2 \
            else np.zeros(shape, np.float32)
    elif callable(initializer):
        val = np.asarray(initializer(tuple(shape)), dtype=np.float32)
    else:
        val = np.full(shape, float(initializer), dtype=np.float32)
    return torch.nn.Parameter(torch.from_numpy(val).to(device), requires_grad=trainable)


def _variable_with_weight_decay(name, shape, wd_value, use_fp16, rng=None, device="cuda"):
    """src/helper_routines.py:49-78 (wd_value is always None at the reference's call sites)."""
    rng = rng or np.random.default_rng(stable_seed(name))
    if wd_value is not None:
        raise NotImplementedError("weight decay is never enabled on the reference's hot path (wd_value=None)")
    return _variable_on_cpu(name, shape, lambda s: variance_scaling(rng, s), use_fp16, device=device)
