"""Evaluation-side consumers of the logits: greedy CTC decoding + character error rate, the path
of src/deepSpeech_test.py:112-136 and :209-213, on the GPU (ds2_ctc_greedy_decode, ds2_edit_distance)."""
from __future__ import annotations

import numpy as np
import torch

from ._lib import check, lib

ALPHABET = "ABCDEFGHIJKLMNOPQRSTUVWXYZ' "      # src/deepSpeech_test.py:28
IX_TO_CHAR = {i: ch for (i, ch) in enumerate(ALPHABET)}


def _stream():
    return torch.cuda.current_stream().cuda_stream


def ctc_greedy_decoder(inputs: torch.Tensor, sequence_length, merge_repeated: bool = True):
    """tf.nn.ctc_greedy_decoder for time-major logits (or log-probabilities) [T,B,K] on the GPU.
    -> (decoded int32 [B,T] CUDA tensor, lengths int32 [B] CUDA tensor); row b holds lengths[b] class ids."""
    if not merge_repeated:
        raise ValueError("ctc_greedy_decoder: only merge_repeated=True (the reference's setting) is built")
    if not inputs.is_cuda:
        raise RuntimeError("ctc_greedy_decoder needs a CUDA tensor (no CPU fallback)")
    x = inputs.contiguous().float()
    T, B, K = x.shape
    lens = torch.as_tensor(np.asarray(sequence_length, dtype=np.int32)).to(x.device) if not torch.is_tensor(sequence_length) \
        else sequence_length.to(device=x.device, dtype=torch.int32)
    out = torch.empty((B, T), dtype=torch.int32, device=x.device)
    out_lens = torch.empty((B,), dtype=torch.int32, device=x.device)
    check(lib.ds2_ctc_greedy_decode(x.data_ptr(), T, B, K, lens.data_ptr(), out.data_ptr(), out_lens.data_ptr(), _stream()),
          "ds2_ctc_greedy_decode")
    return out, out_lens


def edit_distance(a: torch.Tensor, a_lens: torch.Tensor, b: torch.Tensor, b_lens: torch.Tensor) -> torch.Tensor:
    """Levenshtein distance of the rows of two padded int32 CUDA matrices -> int32 [n]."""
    if not (a.is_cuda and b.is_cuda):
        raise RuntimeError("edit_distance needs CUDA tensors (no CPU fallback)")
    a, b = a.contiguous().int(), b.contiguous().int()
    a_lens, b_lens = a_lens.to(a.device).int().contiguous(), b_lens.to(a.device).int().contiguous()
    n = a.shape[0]
    if b.shape[0] != n:
        raise ValueError("edit_distance: %d vs %d sequences" % (n, b.shape[0]))
    out = torch.empty((n,), dtype=torch.int32, device=a.device)
    check(lib.ds2_edit_distance(a.data_ptr(), a.shape[1], a_lens.data_ptr(), b.data_ptr(), b.shape[1], b_lens.data_ptr(), n,
                                max(a.shape[1], b.shape[1]), out.data_ptr(), _stream()), "ds2_edit_distance")
    return out


def pack_label_seqs(seqs, device):
    """list of lists of class ids -> (int32 [n,Lmax], int32 [n]) CUDA tensors."""
    n = len(seqs)
    L = max(1, max((len(s) for s in seqs), default=1))
    m = np.zeros((n, L), np.int32)
    ln = np.zeros((n,), np.int32)
    for i, s in enumerate(seqs):
        m[i, :len(s)] = s
        ln[i] = len(s)
    return torch.from_numpy(m).to(device), torch.from_numpy(ln).to(device)


def char_error_rate(logits: torch.Tensor, seq_lens, label_seqs):
    """Per-utterance CER of the greedy decoding against `label_seqs` (lists of class ids), src/deepSpeech_test.py:112-136.
    -> (list of float, list of decoded strings)."""
    dec, dec_lens = ctc_greedy_decoder(logits, seq_lens)
    lab, lab_lens = pack_label_seqs(label_seqs, logits.device)
    dist = edit_distance(lab, lab_lens, dec, dec_lens).cpu().numpy()
    cer = [float(d) / max(1, len(s)) for d, s in zip(dist, label_seqs)]
    dec_h, lens_h = dec.cpu().numpy(), dec_lens.cpu().numpy()
    texts = ["".join(IX_TO_CHAR.get(int(c), "?") for c in dec_h[i, :lens_h[i]]) for i in range(dec_h.shape[0])]
    return cer, texts
