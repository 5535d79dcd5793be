"""SURVEY 8(f) rank 4 on the CPU: the TFRecord / SequenceExample reader with length bucketing (src/deepSpeech_input.py)
and the checkpoint files (src/deepSpeech_train.py:354-356, :383-398, :511-513)."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch


def _write_split(d, name, lengths, rng):
    from deepspeech_b200 import deepSpeech_input as I
    os.makedirs(os.path.dirname(os.path.join(d, name)), exist_ok=True)
    exs = []
    for n in lengths:
        feats = rng.standard_normal((n, 161)).astype(np.float32)
        labels = rng.integers(0, 28, size=int(rng.integers(1, 9)))
        exs.append((feats, labels))
    I.write_tfrecords(os.path.join(d, name), [I.make_sequence_example(f, l) for f, l in exs])
    return exs


def test_tfrecord_reader_bucketing_and_padding(tmp_path):
    from deepspeech_b200 import deepSpeech_input as I
    assert I.NUM_CLASSES == 29 and I.BUCKET_BOUNDARIES[0] == 100 and I.BUCKET_BOUNDARIES[-1] == 1800
    assert I._crc32c(b"123456789") == 0xE3069283                              # the CRC-32C check value
    rng = np.random.default_rng(0)
    d = str(tmp_path)
    a = _write_split(d, "train-clean-100/train_1.tfrecords", [120, 130, 95, 250, 199, 110], rng)
    b = _write_split(d, "train-clean-100/train_2.tfrecords", [210, 101, 99, 260], rng)
    batches = list(I.inputs("train", d, batch_size=2))
    # buckets: [100,200): 120,130 | 199,110 | then 101 alone (dropped); [0,100): 95,99 ; [200,300): 250,210 | 260 alone
    lens = [tuple(int(x) for x in sl) for _, _, sl in batches]
    assert lens == [(120, 130), (199, 110), (250, 210), (95, 99)]
    feats, (idx, vals, shape), sl = batches[1]
    assert feats.shape == (2, 199, 161) and feats.dtype == np.float32
    assert np.array_equal(feats[0], a[4][0]) and np.array_equal(feats[1, :110], a[5][0]) and not feats[1, 110:].any()
    want = [list(a[4][1]), list(a[5][1])]
    got = [[int(v) for (r, c), v in zip(idx, vals) if r == bb] for bb in range(2)]
    assert got == want and shape[0] == 2 and shape[1] == max(len(w) for w in want) and vals.dtype == np.int32
    # a flipped payload byte is caught by the record checksum
    p = os.path.join(d, "train-clean-100/train_1.tfrecords")
    raw = bytearray(open(p, "rb").read())
    raw[40] ^= 0xFF
    open(p, "wb").write(bytes(raw))
    with pytest.raises(IOError):
        list(I.read_tfrecords(p))
    with pytest.raises(ValueError):
        I.inputs("bogus", d, 2)                                                # unknown split
    os.makedirs(os.path.join(d, "gap", "train-other"))
    open(os.path.join(d, "gap", "train-other", "x.tfrecords"), "wb").close()   # one train* file, but not train-clean-100/train_1
    with pytest.raises(ValueError):
        I.inputs("train", os.path.join(d, "gap"), 2)                           # "Failed to find file" (src/deepSpeech_input.py:88-90)


def test_checkpoint_files_round_trip(tmp_path):
    from deepspeech_b200 import checkpoint as C
    d = str(tmp_path)
    args = SimpleNamespace(num_hidden=1760, num_rnn_layers=7, rnn_type="bi-dir", num_filters=32, initial_lr=1e-4, temporal_stride=2,
                           batch_size=32, train_dir=d)
    C.save_parameters_json(d, args)
    other = SimpleNamespace(num_hidden=1, num_rnn_layers=1, rnn_type="x", num_filters=1, initial_lr=0.0, temporal_stride=0, batch_size=8)
    C.load_parameters_json(d, other)
    assert (other.num_hidden, other.num_rnn_layers, other.rnn_type, other.num_filters, other.initial_lr) == (1760, 7, "bi-dir", 32, 1e-4)
    assert other.batch_size == 8                                               # not one of the restored fields (:118-126)
    m = torch.nn.Linear(5, 3)
    opt = SimpleNamespace(params=list(m.parameters()), m=[torch.randn_like(p) for p in m.parameters()],
                          v=[torch.rand_like(p) for p in m.parameters()], shadow=[torch.randn_like(p) for p in m.parameters()], global_step=30)
    for step in (10, 20, 30):
        C.save(d, m, opt, step)
    assert C.latest_checkpoint(d).endswith("model.ckpt-30.npz")
    m2 = torch.nn.Linear(5, 3)
    opt2 = SimpleNamespace(params=list(m2.parameters()), m=[torch.zeros_like(p) for p in m2.parameters()],
                           v=[torch.zeros_like(p) for p in m2.parameters()], shadow=[torch.zeros_like(p) for p in m2.parameters()], global_step=0)
    assert C.restore(d, m2, opt2) == 30 and opt2.global_step == 30
    for a, b in zip(m.parameters(), m2.parameters()):
        assert torch.equal(a, b)
    for i in range(2):
        assert torch.equal(opt.m[i], opt2.m[i]) and torch.equal(opt.v[i], opt2.v[i]) and torch.equal(opt.shadow[i], opt2.shadow[i])
    m3 = torch.nn.Linear(5, 3)
    C.restore(os.path.join(d, "model.ckpt-20.npz"), m3, use_moving_averages=True)          # what the evaluation script restores
    for s, p in zip(opt.shadow, m3.parameters()):
        assert torch.equal(s, p)
    (tmp_path / "sub").mkdir()
    with pytest.raises(FileNotFoundError):
        C.restore(str(tmp_path / "sub"), m3)                                   # a directory without a checkpoint
