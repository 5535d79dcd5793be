"""`inference` / `loss` with the signatures of the reference's default model module
(src/deepSpeech_NCHW.py:81,204): conv (NCHW semantics: batch_norm2 eps 1e-3, feature = c*75+f) ->
stacked_brnn -> FC -> CTC.  params.wiring selects 'as_written' (the reference verbatim, SURVEY Q1: a
1-live-layer bi-RNN) or 'stacked' (the intended topology; default for throughput runs)."""
from __future__ import annotations

import torch

from . import ctc as _ctc
from .deepSpeech import _model_for
from .model import NUM_CLASSES, get_rnn_seqlen  # noqa: F401


def model(params):
    # engine 'mkldnn_rnn' / 'cudnn_rnn' (src/deepSpeech_NCHW.py:173-176) makes the recurrent layers MkldnnRNNRelu cells with
    # one flat parameter blob each, run through the same sequence kernels (model.DeepSpeech2, mkldnn_rnn_op.py); the NHWC
    # module ignores `engine`, as the reference's does
    return _model_for(params, "NCHW", getattr(params, "wiring", "stacked"))


def inference(sess, feats, seq_lens, params):
    """src/deepSpeech_NCHW.py:81-201.  `sess` accepted and ignored.  Returns logits [T2,B,29]."""
    if feats.shape[0] != params.batch_size:
        raise ValueError("inference: feats batch %d != params.batch_size %d" % (feats.shape[0], params.batch_size))
    return model(params)(feats, seq_lens)


def loss(logits, labels, seq_lens):
    """src/deepSpeech_NCHW.py:204-228 (bare batch mean)."""
    per_utt = _ctc.ctc_loss(labels=labels, inputs=logits, sequence_length=get_rnn_seqlen(seq_lens),
                            preprocess_collapse_repeated=True, time_major=True)
    return torch.mean(per_utt)
