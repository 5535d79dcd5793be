"""`inference` / `loss` / `get_rnn_seqlen` with the signatures of the reference's NHWC model module
(src/deepSpeech.py:38,81,206): conv (NHWC, batch_norm eps 1e-5) -> MultiRNNCell two-stack topology
 -> FC -> CTC."""
from __future__ import annotations

import torch

from . import ctc as _ctc
from .model import NUM_CLASSES, DeepSpeech2, get_rnn_seqlen  # noqa: F401

_MODELS = {}


def _model_for(params, layout, wiring):
    key = (layout, wiring, params.num_filters, params.num_hidden, params.num_rnn_layers, bool(params.use_fp16),
           getattr(params, "rnn_type", "bi-dir"), getattr(params, "cell", "relu_clip"),
           getattr(params, "engine", "mkl") if layout == "NCHW" else "tf")
    m = _MODELS.get(key)
    if m is None:
        m = _MODELS[key] = DeepSpeech2(params.num_filters, params.num_hidden, params.num_rnn_layers, layout=layout,
                                       wiring=wiring, cell=getattr(params, "cell", "relu_clip"), use_bf16=bool(params.use_fp16),
                                       rnn_type=getattr(params, "rnn_type", "bi-dir"), seed=getattr(params, "seed", 1234),
                                       engine=key[-1])
    return m


def model(params):
    """The parameter-holding module behind inference() for this params (the TF graph's variable store)."""
    return _model_for(params, "NHWC", getattr(params, "wiring", "two_stack"))


def inference(session, feats, seq_lens, params):
    """src/deepSpeech.py:81-203.  `session` is accepted and ignored.  feats [B,T,161] f32 (CUDA),
    seq_lens int [B]; params needs .use_fp16 .num_filters .batch_size .num_hidden .num_rnn_layers .rnn_type.
    Returns logits [T2,B,29]."""
    if feats.shape[0] != params.batch_size:
        raise ValueError("inference: feats batch %d != params.batch_size %d" % (feats.shape[0], params.batch_size))
    return model(params)(feats, seq_lens)


def loss(logits, labels, seq_lens):
    """src/deepSpeech.py:206-233: mean over the batch of tf.nn.ctc_loss(labels, float32(logits),
    get_rnn_seqlen(seq_lens), preprocess_collapse_repeated=True, time_major=True)."""
    per_utt = _ctc.ctc_loss(labels=labels, inputs=logits, sequence_length=get_rnn_seqlen(seq_lens),
                            preprocess_collapse_repeated=True, time_major=True)
    return torch.mean(per_utt)
