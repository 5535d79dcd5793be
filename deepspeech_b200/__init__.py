"""deepspeech_b200 - B200-native (sm_100a) implementation of DeepSpeech2's training hot path behind the
operator surface of yxlao/deepSpeech (src/custom_ops.py, src/mkldnn_rnn_op.py, tf.nn.ctc_loss)."""
from . import _lib  # noqa: F401  (fails loudly when the CUDA library is missing)

__all__ = ["_lib"]
