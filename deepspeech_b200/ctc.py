"""ctc_loss with the argument list of tf.nn.ctc_loss as the reference calls it
(src/deepSpeech_NCHW.py:225, src/deepSpeech.py:227)."""
from __future__ import annotations

import torch

from . import ops


def ctc_loss(labels, inputs, sequence_length, preprocess_collapse_repeated=False, ctc_merge_repeated=True,
             ignore_longer_outputs_than_inputs=False, time_major=True):
    """labels: SparseTensor triple (indices [n,2] int64, values [n] int32, dense_shape [2] int64) as built at
    src/deepSpeech_train.py:437; inputs: float logits [T,B,K] (time_major) - softmax is applied inside;
    sequence_length: int32 [B].  Returns the per-utterance negative log-likelihood [B] (float32),
    differentiable w.r.t. inputs.  blank = K-1.  Raises ValueError where TF raises InvalidArgumentError."""
    if not ctc_merge_repeated:
        raise NotImplementedError("ctc_merge_repeated=False is never used by the reference")
    if ignore_longer_outputs_than_inputs:
        raise NotImplementedError("ignore_longer_outputs_than_inputs=True is never used by the reference")
    if not time_major:
        inputs = inputs.transpose(0, 1)
    if inputs.dim() != 3:
        raise ValueError("ctc_loss: inputs must be [max_time, batch, num_classes]")
    indices, values, dense_shape = labels
    K = inputs.shape[2]
    if isinstance(indices, torch.Tensor) and indices.is_cuda and isinstance(values, torch.Tensor) and values.is_cuda:
        # labels already on the device (e.g. deepSpeech_dummy.DeviceDummyData): pre-processing and checks stay there, the
        # step never waits for the host.  A row tf.nn.ctc_loss would raise for - a valid label after a terminator, or fewer
        # frames than labels + repeats - comes back with loss = +inf and a zero gradient row instead of an exception.
        return ops.ctc_loss_sparse_device(inputs, indices, values, dense_shape, sequence_length, preprocess_collapse_repeated)
    lab, ll, reps = ops.ctc_prepare_labels(indices, values, dense_shape, K, preprocess_collapse_repeated)
    return ops.ctc_loss_dense(inputs, lab, ll, reps, sequence_length)
