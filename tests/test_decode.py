"""Greedy CTC decoding + edit distance / CER: oracle known answers on the CPU, kernels vs oracle on the GPU."""
import numpy as np
import pytest
import torch

from oracle import ds2_oracle as O


def test_oracle_greedy_decode_and_edit_distance_known_answers():
    K = 5                                            # blank = 4
    frames = [0, 0, 4, 0, 1, 1, 4, 4, 2, 2, 2, 3]    # A A _ A B B _ _ C C C D  -> A A B C D
    logits = np.full((len(frames), 1, K), -5.0)
    for t, c in enumerate(frames):
        logits[t, 0, c] = 3.0
    assert O.ctc_greedy_decode_np(logits, [len(frames)]) == [[0, 0, 1, 2, 3]]
    assert O.ctc_greedy_decode_np(logits, [4]) == [[0, 0]]                      # sequence length cuts the frames
    assert O.edit_distance_py("KITTEN", "SITTING") == 3
    assert O.edit_distance_py([], [1, 2]) == 2 and O.edit_distance_py([1, 2, 3], [1, 2, 3]) == 0
    assert O.char_error_rates([[1, 2, 3, 4]], [[1, 3, 4]]) == [0.25]


@pytest.mark.gpu
def test_greedy_decode_and_cer_kernels_vs_oracle():
    from deepspeech_b200 import decode
    rng = np.random.default_rng(4)
    T, B, K = 97, 9, 29
    logits = rng.standard_normal((T, B, K)).astype(np.float32)
    logits[:, :, K - 1] += 1.0                       # make blanks and repeats common
    for b in range(B):
        for t in range(1, T):
            if rng.random() < 0.4:
                logits[t, b] = logits[t - 1, b]
    logits[5, 0, 3] = logits[5, 0, 7] = 9.0          # a tie: the lower class wins
    lens = np.asarray([T, 1, 50, 97, 0, 33, 96, 2, 70], np.int32)
    ref = O.ctc_greedy_decode_np(logits, lens)
    x = torch.tensor(logits, device="cuda")
    dec, dec_lens = decode.ctc_greedy_decoder(x, lens)
    dec_h, dl = dec.cpu().numpy(), dec_lens.cpu().numpy()
    for b in range(B):
        assert list(dec_h[b, :dl[b]]) == ref[b], b
    labels = [[int(v) for v in rng.integers(0, 28, size=int(n))] for n in rng.integers(1, 60, size=B)]
    labels[3] = ref[3]                               # an exact match -> CER 0
    cer, texts = decode.char_error_rate(x, lens, labels)
    want = O.char_error_rates(labels, ref)
    assert np.allclose(cer, want) and cer[3] == 0.0
    assert texts[0] == "".join(O.ALPHABET[c] for c in ref[0])
    # longer sequences, empty sides
    a = [[int(v) for v in rng.integers(0, 5, size=n)] for n in (0, 1, 300, 1200, 7)]
    bq = [[int(v) for v in rng.integers(0, 5, size=n)] for n in (4, 0, 310, 1100, 7)]
    A, Al = decode.pack_label_seqs(a, "cuda")
    Bm, Bl = decode.pack_label_seqs(bq, "cuda")
    d = decode.edit_distance(A, Al, Bm, Bl).cpu().numpy()
    assert list(d) == [O.edit_distance_py(p, q) for p, q in zip(a, bq)]
    with pytest.raises(RuntimeError):
        decode.ctc_greedy_decoder(torch.zeros(3, 2, 29), [3, 3])


@pytest.mark.gpu
def test_label_transform_on_the_device_matches_the_host_form():
    """ds2_ctc_labels_from_sparse against ops.ctc_prepare_labels (itself checked against the oracle in test_host.py)."""
    from deepspeech_b200 import ops
    rng = np.random.default_rng(6)
    B, L, K = 7, 23, 29
    dense = rng.integers(0, 28, size=(B, L))
    dense[rng.random((B, L)) < 0.35] = 0                       # zeros are dropped by the sparse conversion (dummy.py:89-98)
    for b in range(B):
        for j in range(1, L):
            if rng.random() < 0.3:
                dense[b, j] = dense[b, j - 1]                      # adjacent repeats: collapsed when asked
    dense[2, 9:] = 0
    dense[2, 9] = 28                                              # a terminator with nothing valid after it
    dense[5] = 0                                                  # an empty row
    idx = np.argwhere(dense != 0).astype(np.int64)
    vals = dense[dense != 0].astype(np.int32)
    for collapse in (True, False):
        lab_h, len_h, rep_h = ops.ctc_prepare_labels(idx, vals, dense.shape, K, collapse)
        lab_d, len_d, rep_d, st = ops.ctc_prepare_labels_device(torch.tensor(idx, device="cuda"), torch.tensor(vals, device="cuda"),
                                                                dense.shape, K, collapse)
        assert st.cpu().numpy().tolist() == [0] * B
        assert len_d.cpu().numpy().tolist() == len_h.tolist() and rep_d.cpu().numpy().tolist() == rep_h.tolist()
        ld = lab_d.cpu().numpy()
        for b in range(B):
            assert ld[b, :len_h[b]].tolist() == lab_h[b, :len_h[b]].tolist() and not ld[b, len_h[b]:].any()
    # a valid label after a terminator: the host form raises, the device form flags the row
    bad_idx = np.asarray([[0, 0], [0, 1], [1, 0]], np.int64)
    bad_vals = np.asarray([28, 3, 5], np.int32)
    with pytest.raises(ValueError):
        ops.ctc_prepare_labels(bad_idx, bad_vals, (2, 2), K, True)
    _, _, _, st = ops.ctc_prepare_labels_device(torch.tensor(bad_idx, device="cuda"), torch.tensor(bad_vals, device="cuda"), (2, 2), K, True)
    assert st.cpu().numpy().tolist() == [1, 0]
