"""Checkpoints of the training loop, the on-disk formats either side of the hot path:

* `deepSpeech_parameters.json` - the run's hyper-parameters, written once into train_dir and re-read on resume
  (src/deepSpeech_train.py:511-513 writes `json.dumps(vars(ARGS))`, :114-127 reads it back);
* `model.ckpt-<step>.npz` + the text file `checkpoint` - what `tf.train.Saver(tf.global_variables(), max_to_keep=100)`
  keeps (src/deepSpeech_train.py:471, saved every 10 steps at :354-356, restored through `get_checkpoint_state` at
  :383-398): every trainable variable, the optimizer's Adam slots m / v, the moving-average shadow of every trainable
  (`<name>/ExponentialMovingAverage`, restored INSTEAD of the weights by the evaluation script,
  src/deepSpeech_test.py:218-220) and the global step.

The container is numpy's .npz (TensorFlow's own checkpoint format needs TensorFlow): one array per variable under the
variable's name.  Values are fp32 bit patterns: the restored state equals the saved one bit for bit and the next step's loss
is identical (tests/test_gpu_checkpoint.py; its gradients are atomic sums, reproducible to fp32 round-off only).
Host-side code: tensors cross the PCIe bus once per save / restore, nothing here is on the timed path."""
from __future__ import annotations

import json
import os
import re
from types import SimpleNamespace
from typing import Optional

import numpy as np
import torch

PARAMS_FILE = "deepSpeech_parameters.json"      # src/deepSpeech_train.py:511
STATE_FILE = "checkpoint"                        # tf.train.get_checkpoint_state's index file
EMA_SUFFIX = "/ExponentialMovingAverage"
MAX_TO_KEEP = 100                                # src/deepSpeech_train.py:471


def save_parameters_json(train_dir: str, args) -> str:
    """src/deepSpeech_train.py:511-513."""
    os.makedirs(train_dir, exist_ok=True)
    d = dict(vars(args)) if not isinstance(args, dict) else dict(args)
    path = os.path.join(train_dir, PARAMS_FILE)
    with open(path, "w") as f:
        f.write(json.dumps({k: v for k, v in d.items() if isinstance(v, (int, float, str, bool, type(None), list))}))
    return path


def load_parameters_json(checkpoint_dir: str, args=None):
    """src/deepSpeech_train.py:114-127: on resume the model-defining hyper-parameters come from the checkpoint directory,
    not from the command line.  Returns a namespace; with `args` given, its fields are overwritten in place."""
    with open(os.path.join(checkpoint_dir, PARAMS_FILE)) as f:
        d = json.load(f)
    if args is None:
        return SimpleNamespace(**d)
    for k in ("num_hidden", "num_rnn_layers", "rnn_type", "num_filters", "initial_lr", "temporal_stride"):   # :118-126
        if k in d:
            setattr(args, k, d[k])
    return args


def _ckpt_name(step: int) -> str:
    return "model.ckpt-%d" % int(step)


def _read_state(train_dir: str):
    path = os.path.join(train_dir, STATE_FILE)
    if not os.path.exists(path):
        return None, []
    latest, allp = None, []
    for line in open(path):
        m = re.match(r'\s*(model_checkpoint_path|all_model_checkpoint_paths):\s*"(.*)"', line)
        if m:
            if m.group(1) == "model_checkpoint_path":
                latest = m.group(2)
            else:
                allp.append(m.group(2))
    return latest, allp


def _write_state(train_dir: str, latest: str, allp):
    with open(os.path.join(train_dir, STATE_FILE), "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % latest)
        for p in allp:
            f.write('all_model_checkpoint_paths: "%s"\n' % p)


def save(train_dir: str, model: torch.nn.Module, optimizer=None, step: Optional[int] = None) -> str:
    """saver.save(sess, os.path.join(train_dir, 'model.ckpt'), global_step=step): src/deepSpeech_train.py:354-356."""
    os.makedirs(train_dir, exist_ok=True)
    if step is None:
        step = optimizer.global_step if optimizer is not None else 0
    arrays = {"global_step": np.asarray(int(step), np.int64)}
    names = {}
    for name, p in model.named_parameters():
        arrays[name] = p.detach().cpu().numpy()
        names[id(p)] = name
    if optimizer is not None:
        for i, p in enumerate(optimizer.params):
            n = names[id(p)]
            arrays[n + "/Adam"] = optimizer.m[i].cpu().numpy()            # TF's slot names
            arrays[n + "/Adam_1"] = optimizer.v[i].cpu().numpy()
            if optimizer.shadow is not None:
                arrays[n + EMA_SUFFIX] = optimizer.shadow[i].cpu().numpy()
    base = _ckpt_name(step)
    tmp = os.path.join(train_dir, base + ".tmp.npz")
    np.savez(tmp, **arrays)
    os.replace(tmp, os.path.join(train_dir, base + ".npz"))
    _, allp = _read_state(train_dir)
    allp = [p for p in allp if p != base] + [base]
    while len(allp) > MAX_TO_KEEP:                                         # max_to_keep: the oldest file goes
        old = allp.pop(0)
        try:
            os.remove(os.path.join(train_dir, old + ".npz"))
        except OSError:
            pass
    _write_state(train_dir, base, allp)
    return os.path.join(train_dir, base + ".npz")


def latest_checkpoint(checkpoint_dir: str) -> Optional[str]:
    """tf.train.get_checkpoint_state(dir).model_checkpoint_path (src/deepSpeech_train.py:386-390)."""
    latest, _ = _read_state(checkpoint_dir)
    if latest is None:
        return None
    path = os.path.join(checkpoint_dir, os.path.basename(latest) + ".npz")
    return path if os.path.exists(path) else None


def restore(path_or_dir: str, model: torch.nn.Module, optimizer=None, use_moving_averages: bool = False) -> int:
    """Restores weights (or, with use_moving_averages, their EMA shadows - src/deepSpeech_test.py:218-220), the Adam slots
    and the step counter.  -> the global step of the checkpoint (the reference parses it from the file name, :392-394).
    Raises FileNotFoundError like the reference's `return` path would leave the session uninitialised."""
    path = path_or_dir
    if os.path.isdir(path_or_dir):
        path = latest_checkpoint(path_or_dir)
        if path is None:
            raise FileNotFoundError("no checkpoint found in %s" % path_or_dir)
    with np.load(path) as z:
        step = int(z["global_step"])
        with torch.no_grad():
            for name, p in model.named_parameters():
                key = name + EMA_SUFFIX if use_moving_averages else name
                if key not in z.files:
                    raise KeyError("checkpoint %s has no variable %s" % (path, key))
                a = z[key]
                if tuple(a.shape) != tuple(p.shape):
                    raise ValueError("checkpoint variable %s has shape %s, the model's is %s" % (key, a.shape, tuple(p.shape)))
                p.copy_(torch.from_numpy(a))
        if optimizer is not None:
            names = {id(p): n for n, p in model.named_parameters()}
            for i, p in enumerate(optimizer.params):
                n = names[id(p)]
                optimizer.m[i].copy_(torch.from_numpy(z[n + "/Adam"]))
                optimizer.v[i].copy_(torch.from_numpy(z[n + "/Adam_1"]))
                if optimizer.shadow is not None and n + EMA_SUFFIX in z.files:
                    optimizer.shadow[i].copy_(torch.from_numpy(z[n + EMA_SUFFIX]))
            optimizer.global_step = step
    try:
        from . import ops
        ops.invalidate_weight_cache()
    except Exception:
        pass
    return step
