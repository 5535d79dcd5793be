"""Synthetic batch generator of the reference's `dummy=1` mode, restated (src/deepSpeech_dummy.py).

15 length buckets; features are contiguous slices of a pre-drawn randn pool; ONE random label sequence
per batch shared by every utterance, last label = NUM_CLASSES-1 (TF's blank index), zeros dropped by
`_dense_to_sparse`.  Host numpy, seeded (the reference is unseeded)."""
from __future__ import annotations

import numpy as np

utt_lengths = [100, 200, 100, 400, 500, 600, 700, 800, 900, 1000, 1100, 1200, 1300, 1400, 1500]   # dummy.py:9
counts = [30, 3, 3, 3, 14, 13, 9, 8, 5, 4, 3, 2, 2, 2, 1]                                            # dummy.py:10
label_lengths = [7, 17, 7, 48, 62, 78, 93, 107, 120, 134, 148, 163, 178, 193, 209]                   # dummy.py:11
freq_bins = 161
scale_factor = 10
extra = 1000
NUM_CLASSES = 29
NUM_PER_EPOCH_FOR_TRAIN = sum(counts) * scale_factor


class DummyData(object):
    """State machine of dummy.py:32-87 (`_init_data` / `_next`), one instance per process."""

    def __init__(self, batch_size, seed=1234):
        self.batch_size = batch_size
        self.rng = np.random.default_rng(seed)
        line = batch_size * (utt_lengths[-1] + extra)
        self.pool = self.rng.standard_normal((line, freq_bins), dtype=np.float32)      # dummy.py:47
        self.utter_counts = [x * scale_factor * batch_size for x in counts]               # dummy.py:43
        self.current = 0

    def _advance(self):
        """dummy.py:58-75: the bucket of this step; buckets are visited in order, counts[i] * 10 batches each."""
        bs = self.batch_size
        if self.current >= len(self.utter_counts):                                        # dummy.py:58-61
            self.current = 0
            self.utter_counts = [x * scale_factor * bs for x in counts]
        cur = self.current
        if self.utter_counts[cur] > bs:
            self.utter_counts[cur] -= bs
        else:
            self.utter_counts[cur] = 0
            self.current += 1
        return cur

    def _next(self, bucket=None):
        bs = self.batch_size
        cur = self._advance() if bucket is None else bucket
        utt_length, label_length = utt_lengths[cur], label_lengths[cur]
        start = int(self.rng.integers(0, extra + bs * (utt_lengths[-1] - utt_length)))   # dummy.py:77
        feat = self.pool[start:start + utt_length * bs]
        label = self.rng.integers(0, NUM_CLASSES - 1, size=label_length)                  # randint(0, 27)
        label[label_length - 1] = NUM_CLASSES - 1                                         # dummy.py:84
        return utt_length, feat, label

    def inputs(self, bucket=None):
        """-> feats [B,T,161] f32, idx [n,2] i64, vals [n] i32, dense_shape [2] i64, seq_lens [B] i32
        (src/deepSpeech_dummy.py:101-122)."""
        bs = self.batch_size
        utt_length, feat, label = self._next(bucket)
        feats = feat.reshape(bs, utt_length, freq_bins)
        nz = np.nonzero(label != 0)[0]                                                    # _dense_to_sparse, :89-98
        idx = np.stack([np.repeat(np.arange(bs), nz.size), np.tile(nz, bs)], 1).astype(np.int64)
        vals = np.tile(label[nz], bs).astype(np.int32)
        return (np.ascontiguousarray(feats, dtype=np.float32), idx, vals, np.asarray([bs, len(label)], dtype=np.int64),
                np.full(bs, utt_length, dtype=np.int32))


class DeviceDummyData(DummyData):
    """The same state machine with the feature pool RESIDENT ON THE GPU: a batch's features are a
    zero-copy slice of the device pool, its SparseTensor labels are built by a kernel from the <= 209 non-zero entries of
    the shared label row.  Same seed -> the same buckets, offsets and labels as the host generator, bit for bit; no
    per-step host loop over B x L entries (src/deepSpeech_dummy.py:89-98) and no 31 MB host -> device copy."""

    def __init__(self, batch_size, seed=1234, device="cuda"):
        super().__init__(batch_size, seed)
        import torch
        self.device = torch.device(device)
        self.pool_dev = torch.from_numpy(self.pool).to(self.device)

    def inputs(self, bucket=None):
        """-> feats [B,T,161] f32 (CUDA, a view of the pool), idx [n,2] i64 (CUDA), vals [n] i32 (CUDA),
        dense_shape (host, 2 ints), seq_lens [B] i32 (host numpy, as the reference feeds them)."""
        import torch
        from . import ops
        bs = self.batch_size
        if bucket is None:
            cur = self._advance()
        else:
            cur = bucket
        utt_length, label_length = utt_lengths[cur], label_lengths[cur]
        start = int(self.rng.integers(0, extra + bs * (utt_lengths[-1] - utt_length)))
        label = self.rng.integers(0, NUM_CLASSES - 1, size=label_length)
        label[label_length - 1] = NUM_CLASSES - 1
        feats = self.pool_dev[start:start + utt_length * bs].view(bs, utt_length, freq_bins)
        nz = np.nonzero(label != 0)[0].astype(np.int32)
        host = torch.from_numpy(np.concatenate([nz, label[nz].astype(np.int32)])).to(self.device, non_blocking=True)
        n = int(nz.size)
        idx = torch.empty((bs * n, 2), dtype=torch.int64, device=self.device)
        vals = torch.empty((bs * n,), dtype=torch.int32, device=self.device)
        ops.check(ops.lib.ds2_dummy_sparse_labels(ops._p(host[:n]), ops._p(host[n:]), n, bs, ops._p(idx), ops._p(vals), ops._stream()),
                  "ds2_dummy_sparse_labels")
        return feats, idx, vals, np.asarray([bs, len(label)], dtype=np.int64), np.full(bs, utt_length, dtype=np.int32)


_STATE = {}


def inputs(batch_size, seed=1234):
    """Module-level entry with the reference's signature (dummy.py:101)."""
    st = _STATE.get(batch_size)
    if st is None:
        st = _STATE[batch_size] = DummyData(batch_size, seed)
    return st.inputs()
