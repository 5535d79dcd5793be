"""Real-data reader of the reference, restated without TensorFlow (src/deepSpeech_input.py:20-97):
TFRecord files of `tf.train.SequenceExample`s (context: seq_len int64, labels int64 list; feature list `feats`: one
161-float frame per step - written by src/preprocess_LibriSpeech.py:45-85) -> length-bucketed, zero-padded batches
(`tf.contrib.training.bucket_by_sequence_length`, boundaries 100..1800, dynamic_pad, no smaller final batch).

Host-side Python + numpy: the wire formats are parsed by hand (TFRecord framing with masked CRC-32C, protobuf varints),
because neither TensorFlow nor its example protos exist in this image.  Nothing here is on the timed path."""
from __future__ import annotations

import bisect
import glob
import os
import struct
from typing import Iterable, Iterator, List, Tuple

import numpy as np

ALPHABET = "ABCDEFGHIJKLMNOPQRSTUVWXYZ' "          # src/deepSpeech_input.py:13
NUM_CLASSES = len(ALPHABET) + 1                    # :14, additional class for the blank
NUM_PER_EPOCH_FOR_TRAIN = 28535                    # :15
NUM_PER_EPOCH_FOR_EVAL = 2703
NUM_PER_EPOCH_FOR_TEST = 2620
FREQ_BINS = 161
BUCKET_BOUNDARIES = list(range(100, 1900, 100))    # :57

# ----------------------------------------------------------------------------- CRC-32C (Castagnoli), TFRecord's checksum
_CRC_TABLE = []
for _i in range(256):
    _c = _i
    for _ in range(8):
        _c = (_c >> 1) ^ 0x82F63B78 if _c & 1 else _c >> 1
    _CRC_TABLE.append(_c)


def _crc32c(data: bytes) -> int:
    c = 0xFFFFFFFF
    for b in data:
        c = _CRC_TABLE[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def _masked_crc(data: bytes) -> int:
    c = _crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


# ----------------------------------------------------------------------------- TFRecord framing
def read_tfrecords(path: str, verify_crc: bool = True) -> Iterator[bytes]:
    """uint64 length | uint32 masked crc(length) | payload | uint32 masked crc(payload), little endian."""
    with open(path, "rb") as f:
        while True:
            head = f.read(12)
            if not head:
                return
            if len(head) < 12:
                raise IOError("%s: truncated record header" % path)
            (n,), (hcrc,) = struct.unpack("<Q", head[:8]), struct.unpack("<I", head[8:])
            if verify_crc and _masked_crc(head[:8]) != hcrc:
                raise IOError("%s: corrupt record length" % path)
            data = f.read(n)
            tail = f.read(4)
            if len(data) < n or len(tail) < 4:
                raise IOError("%s: truncated record" % path)
            if verify_crc and _masked_crc(data) != struct.unpack("<I", tail)[0]:
                raise IOError("%s: corrupt record payload" % path)
            yield data


def write_tfrecords(path: str, payloads: Iterable[bytes]) -> None:
    with open(path, "wb") as f:
        for data in payloads:
            head = struct.pack("<Q", len(data))
            f.write(head + struct.pack("<I", _masked_crc(head)) + data + struct.pack("<I", _masked_crc(data)))


# ----------------------------------------------------------------------------- protobuf wire format (the subset used)
def _varint(buf: bytes, i: int) -> Tuple[int, int]:
    v, shift = 0, 0
    while True:
        b = buf[i]
        i += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, i
        shift += 7


def _fields(buf: bytes) -> Iterator[Tuple[int, int, object]]:
    """(field number, wire type, value) of one message; length-delimited values come back as bytes."""
    i, n = 0, len(buf)
    while i < n:
        key, i = _varint(buf, i)
        fn, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _varint(buf, i)
        elif wt == 2:
            ln, i = _varint(buf, i)
            v = buf[i:i + ln]
            i += ln
        elif wt == 5:
            v = buf[i:i + 4]
            i += 4
        elif wt == 1:
            v = buf[i:i + 8]
            i += 8
        else:
            raise ValueError("unsupported protobuf wire type %d" % wt)
        yield fn, wt, v


def _enc_varint(v: int) -> bytes:
    out = bytearray()
    v &= (1 << 64) - 1
    while True:
        b = v & 0x7F
        v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def _ld(fn: int, payload: bytes) -> bytes:
    return _enc_varint((fn << 3) | 2) + _enc_varint(len(payload)) + payload


def _parse_feature(buf: bytes):
    """tf.train.Feature: bytes_list = 1, float_list = 2, int64_list = 3 (each: repeated value = 1, packed or not)."""
    for fn, _, v in _fields(buf):
        if fn == 2:
            vals = []
            for f2, wt, x in _fields(v):
                if f2 == 1:
                    vals.append(np.frombuffer(x, dtype="<f4"))          # packed (wt 2) or a single fixed32 (wt 5)
            return np.concatenate(vals) if vals else np.zeros(0, np.float32)
        if fn == 3:
            vals: List[int] = []
            for f2, wt, x in _fields(v):
                if f2 != 1:
                    continue
                if wt == 0:
                    vals.append(x)
                else:
                    j = 0
                    while j < len(x):
                        y, j = _varint(x, j)
                        vals.append(y)
            a = np.asarray(vals, dtype=np.uint64).astype(np.int64)
            return a
        if fn == 1:
            return [x for f2, _, x in _fields(v) if f2 == 1]
    return None


def parse_sequence_example(buf: bytes):
    """tf.parse_single_sequence_example with the reference's feature spec (src/deepSpeech_input.py:36-50):
    -> (seq_len int, labels int64 [n], feats float32 [T,161])."""
    seq_len, labels, frames = None, np.zeros(0, np.int64), []
    for fn, _, v in _fields(buf):
        if fn == 1:                                                    # context: Features{ map<string, Feature> feature = 1 }
            for f2, _, entry in _fields(v):
                key, feat = None, None
                for f3, _, x in _fields(entry):
                    if f3 == 1:
                        key = x.decode()
                    elif f3 == 2:
                        feat = _parse_feature(x)
                if key == "seq_len":
                    seq_len = int(feat[0])
                elif key == "labels":
                    labels = np.asarray(feat, np.int64)
        elif fn == 2:                                                  # feature_lists: map<string, FeatureList>
            for f2, _, entry in _fields(v):
                key, flist = None, None
                for f3, _, x in _fields(entry):
                    if f3 == 1:
                        key = x.decode()
                    elif f3 == 2:
                        flist = x
                if key == "feats" and flist is not None:
                    frames = [_parse_feature(x) for f4, _, x in _fields(flist) if f4 == 1]
    feats = np.stack(frames).astype(np.float32) if frames else np.zeros((0, FREQ_BINS), np.float32)
    if feats.shape[1:] != (FREQ_BINS,):
        raise ValueError("feats frames must be %d-dimensional, got %s" % (FREQ_BINS, feats.shape))
    if seq_len is None:
        raise ValueError("SequenceExample without context feature seq_len")
    return seq_len, labels, feats


def make_sequence_example(feats: np.ndarray, labels) -> bytes:
    """The writer's side (src/preprocess_LibriSpeech.py:45-85 `make_example`): context seq_len + labels, feature list
    feats with one FloatList of 161 values per frame.  Used to build fixtures; byte-compatible with TF's serialisation."""
    feats = np.asarray(feats, dtype="<f4")

    def feature_int64(vals):
        packed = b"".join(_enc_varint(int(v)) for v in vals)
        return _ld(3, _ld(1, packed))

    def feature_float(row):
        return _ld(2, _ld(1, row.tobytes()))

    def map_entry(key, value):
        return _ld(1, key.encode()) + _ld(2, value)
    context = _ld(1, map_entry("seq_len", feature_int64([feats.shape[0]]))) + _ld(1, map_entry("labels", feature_int64(labels)))
    flist = b"".join(_ld(1, feature_float(row)) for row in feats)
    feature_lists = _ld(1, map_entry("feats", flist))
    return _ld(1, context) + _ld(2, feature_lists)


# ----------------------------------------------------------------------------- bucketing
def bucket_by_sequence_length(examples: Iterable[Tuple[int, np.ndarray, np.ndarray]], batch_size: int,
                              bucket_boundaries=BUCKET_BOUNDARIES, allow_smaller_final_batch: bool = False):
    """tf.contrib.training.bucket_by_sequence_length(..., dynamic_pad=True) as called at src/deepSpeech_input.py:53-60: an
    example of length n goes to bucket bisect_right(boundaries, n); a bucket that has batch_size examples yields
    (feats [B,Tmax,161] zero-padded, labels as a SparseTensor triple, seq_len int32 [B]).  The reference drops the
    examples left in partly filled buckets at the end (allow_smaller_final_batch=False)."""
    buckets: List[List] = [[] for _ in range(len(bucket_boundaries) + 1)]
    for ex in examples:
        q = buckets[bisect.bisect_right(bucket_boundaries, ex[0])]
        q.append(ex)
        if len(q) == batch_size:
            yield _pad_batch(q)
            del q[:]
    if allow_smaller_final_batch:
        for q in buckets:
            if q:
                yield _pad_batch(q)


def _pad_batch(items):
    B = len(items)
    Tmax = max(f.shape[0] for _, _, f in items)
    Lmax = max([len(l) for _, l, _ in items] + [1])
    feats = np.zeros((B, Tmax, FREQ_BINS), np.float32)
    idx, vals = [], []
    for b, (n, lab, f) in enumerate(items):
        feats[b, :f.shape[0]] = f
        for j, v in enumerate(lab):
            idx.append((b, j))
            vals.append(int(v))
    return (feats, (np.asarray(idx, np.int64).reshape(-1, 2), np.asarray(vals, np.int32), np.asarray([B, Lmax], np.int64)),
            np.asarray([n for n, _, _ in items], np.int32))


def inputs(eval_data: str, data_dir: str, batch_size: int, shuffle: bool = False, seed: int = 0, epochs: int = 1):
    """src/deepSpeech_input.py:65-97: file lists per split (train: train-clean-100/train_<i>.tfrecords, i = 1..N; val:
    dev*/*.tfrecords; test: test*/*.tfrecords), missing files raise ValueError, then bucketed batches.  The reference's
    filename queue cycles for ever; here `epochs` bounds it (None = for ever)."""
    if eval_data == "train":
        num_files = len(glob.glob(os.path.join(data_dir, "train*/*.tfrecords")))
        filenames = [os.path.join(data_dir, "train-clean-100/train_" + str(i) + ".tfrecords") for i in range(1, num_files + 1)]
    elif eval_data == "val":
        filenames = sorted(glob.glob(os.path.join(data_dir, "dev*/*.tfrecords")))
    elif eval_data == "test":
        filenames = sorted(glob.glob(os.path.join(data_dir, "test*/*.tfrecords")))
    else:
        raise ValueError("eval_data must be 'train', 'val' or 'test'")
    for f in filenames:
        if not os.path.exists(f):
            raise ValueError("Failed to find file: " + f)
    rng = np.random.default_rng(seed)

    def examples():
        ep = 0
        while epochs is None or ep < epochs:
            order = list(filenames)
            if shuffle:
                rng.shuffle(order)
            for f in order:
                for rec in read_tfrecords(f):
                    yield parse_sequence_example(rec)
            ep += 1
    return bucket_by_sequence_length(examples(), batch_size)
