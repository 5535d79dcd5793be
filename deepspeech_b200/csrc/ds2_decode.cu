// Evaluation-side consumers of the logits: greedy CTC decoding and the edit distance behind the
// character error rate.  Reference: tf.nn.ctc_greedy_decoder(tf.nn.log_softmax(logits), get_rnn_seqlen(seq_lens)) at
// src/deepSpeech_test.py:209-213 (merge_repeated=True: collapse runs of the per-frame argmax, then drop the blank
// K-1; log_softmax does not change the argmax and is skipped) and Levenshtein.distance(label, prediction) / len(label)
// at src/deepSpeech_test.py:126-130.
#include "common.cuh"

namespace ds2 {

// one warp per utterance; lane k holds class k (K <= 32), ties go to the lowest class index (Eigen argmax)
__global__ void __launch_bounds__(128)
ctc_greedy_kernel(const float* __restrict__ logits, int T, int B, int K, const int* __restrict__ seq_lens, int* __restrict__ out,
                  int* __restrict__ out_lens) {
  const int b = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (b >= B) return;
  int n = seq_lens[b];
  n = n < 0 ? 0 : (n > T ? T : n);
  int prev = -1, len = 0;
  const int blank = K - 1;
  for (int t = 0; t < n; ++t) {
    float v = lane < K ? logits[((size_t)t * B + b) * K + lane] : -INFINITY;
    int idx = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, v, o);
      const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
      if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    if (idx != prev && idx != blank) {
      if (lane == 0) out[(size_t)b * T + len] = idx;
      ++len;
    }
    prev = idx;
  }
  if (lane == 0) out_lens[b] = len;
}

// one block per pair of sequences; anti-diagonal wavefront over the (la+1) x (lb+1) table, three diagonals in smem
__global__ void __launch_bounds__(256)
edit_distance_kernel(const int* __restrict__ a, int lda, const int* __restrict__ a_lens, const int* __restrict__ b, int ldb,
                     const int* __restrict__ b_lens, int* __restrict__ out, int maxlen) {
  extern __shared__ int ed_smem[];
  const int u = blockIdx.x;
  const int la = a_lens[u], lb = b_lens[u];
  const int* sa = a + (size_t)u * lda;
  const int* sb = b + (size_t)u * ldb;
  int* d0 = ed_smem;                       // diagonal d-2, indexed by i
  int* d1 = d0 + (maxlen + 1);             // diagonal d-1
  int* d2 = d1 + (maxlen + 1);             // diagonal d
  if (la == 0 || lb == 0) {
    if (threadIdx.x == 0) out[u] = la + lb;
    return;
  }
  for (int d = 0; d <= la + lb; ++d) {
    const int i_lo = d > lb ? d - lb : 0, i_hi = d < la ? d : la;
    for (int i = i_lo + threadIdx.x; i <= i_hi; i += blockDim.x) {
      const int j = d - i;
      int v;
      if (i == 0) v = j;
      else if (j == 0) v = i;
      else {
        const int sub = d0[i - 1] + (sa[i - 1] != sb[j - 1]);
        const int del = d1[i - 1] + 1;      // (i-1, j)
        const int ins = d1[i] + 1;          // (i, j-1)
        v = min(sub, min(del, ins));
      }
      d2[i] = v;
    }
    __syncthreads();
    int* t = d0; d0 = d1; d1 = d2; d2 = t;
  }
  if (threadIdx.x == 0) out[u] = d1[la];
}

}  // namespace ds2

using namespace ds2;

extern "C" int ds2_ctc_greedy_decode(const float* logits, int T, int B, int K, const int* seq_lens, int* decoded, int* decoded_lens,
                                     void* stream) {
  DS2_REQUIRE(logits && seq_lens && decoded && decoded_lens, DS2_ERR_BAD_ARG, "ds2_ctc_greedy_decode: null pointer");
  DS2_REQUIRE(T > 0 && B > 0 && K >= 2 && K <= 32, DS2_ERR_BAD_SHAPE, "ds2_ctc_greedy_decode: bad shape T=%d B=%d K=%d (K <= 32)", T, B, K);
  ctc_greedy_kernel<<<(B + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(logits, T, B, K, seq_lens, decoded, decoded_lens);
  DS2_LAUNCH_CHECK("ctc_greedy_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_edit_distance(const int* a, int lda, const int* a_lens, const int* b, int ldb, const int* b_lens, int n, int maxlen,
                                 int* out, void* stream) {
  DS2_REQUIRE(a && a_lens && b && b_lens && out, DS2_ERR_BAD_ARG, "ds2_edit_distance: null pointer");
  DS2_REQUIRE(n > 0 && maxlen >= 0 && maxlen <= 4000 && lda >= 0 && ldb >= 0, DS2_ERR_BAD_SHAPE,
              "ds2_edit_distance: bad shape n=%d maxlen=%d (sequences of at most 4000 symbols)", n, maxlen);
  edit_distance_kernel<<<n, 256, 3 * (maxlen + 1) * sizeof(int), static_cast<cudaStream_t>(stream)>>>(a, lda, a_lens, b, ldb, b_lens, out,
                                                                                                    maxlen);
  DS2_LAUNCH_CHECK("edit_distance_kernel");
  count_launch();
  return DS2_OK;
}

// ------------------------------------------------------------------ label pre-processing on the device
// tf.nn.ctc_loss's handling of the SparseTensor labels built at src/deepSpeech_train.py:437 (values produced by the
// host double loop of src/deepSpeech_dummy.py:89-98): per batch row, in index order, collapse adjacent repeats when
// asked, stop at the first label >= num_classes-1; a valid label after such a terminator is an error (status 1).
// indices [n,2] int64 sorted row-major (the SparseTensor convention), values [n] int32.
namespace ds2 {
__global__ void ctc_labels_from_sparse_kernel(const long long* __restrict__ indices, const int* __restrict__ values, int n, int B, int Lcap,
                                              int num_classes, int collapse, int* __restrict__ labels, int* __restrict__ label_lens,
                                              int* __restrict__ repeats, int* __restrict__ status) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  int lo = 0, hi = n;                      // first entry of row b
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (indices[2 * (long long)mid] < b) lo = mid + 1; else hi = mid;
  }
  int len = 0, rep = 0, st = 0;
  long long prev_col = -1;
  int prev_v = -1, last = -1;
  bool have_prev = false, finished = false;
  for (int i = lo; i < n && indices[2 * (long long)i] == b; ++i) {
    const long long col = indices[2 * (long long)i + 1];
    if (col < prev_col) st |= 2;           // not sorted within the row
    prev_col = col;
    const int v = values[i];
    if (!have_prev || !collapse || v != prev_v) {
      if (v >= num_classes - 1 || v < 0) {
        finished = true;
      } else if (finished) {
        st |= 1;
      } else if (len < Lcap) {
        labels[(long long)b * Lcap + len] = v;
        if (len > 0 && v == last) ++rep;
        last = v;
        ++len;
      } else {
        st |= 4;                           // more labels than the dense shape allows
      }
    }
    prev_v = v;
    have_prev = true;
  }
  for (int j = len; j < Lcap; ++j) labels[(long long)b * Lcap + j] = 0;
  label_lens[b] = len;
  repeats[b] = rep;
  status[b] = st;
}
}  // namespace ds2

extern "C" int ds2_ctc_labels_from_sparse(const long long* indices, const int* values, int n, int B, int Lcap, int num_classes,
                                          int collapse_repeated, int* labels, int* label_lens, int* repeats, int* status,
                                          void* stream) {
  DS2_REQUIRE(labels && label_lens && repeats && status && (n == 0 || (indices && values)), DS2_ERR_BAD_ARG,
              "ds2_ctc_labels_from_sparse: null pointer");
  DS2_REQUIRE(n >= 0 && B > 0 && Lcap > 0 && num_classes >= 2, DS2_ERR_BAD_SHAPE, "ds2_ctc_labels_from_sparse: bad shape");
  ctc_labels_from_sparse_kernel<<<(B + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream)>>>(indices, values, n, B, Lcap, num_classes,
                                                                                             collapse_repeated, labels, label_lens,
                                                                                             repeats, status);
  DS2_LAUNCH_CHECK("ctc_labels_from_sparse_kernel");
  count_launch();
  return DS2_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Dummy-batch labels on the device (src/deepSpeech_dummy.py:89-98 `_dense_to_sparse` + :115-118): every utterance of a
// batch carries the SAME label row; the zero labels are dropped.  nz_pos / nz_val are that row's non-zero positions and
// values (<= 209 entries); the SparseTensor triple of src/deepSpeech_train.py:437 comes back as indices [B*nnz, 2]
// (row-major sorted) and values [B*nnz].
namespace ds2 {
__global__ void __launch_bounds__(256) dummy_sparse_kernel(const int* __restrict__ nz_pos, const int* __restrict__ nz_val, int nnz, int B,
                                                           long long* __restrict__ indices, int* __restrict__ values) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * nnz) return;
  const int b = i / nnz, j = i - b * nnz;
  indices[2 * (size_t)i] = b;
  indices[2 * (size_t)i + 1] = nz_pos[j];
  values[i] = nz_val[j];
}
}  // namespace ds2

extern "C" int ds2_dummy_sparse_labels(const int* nz_pos, const int* nz_val, int nnz, int B, long long* indices, int* values, void* stream) {
  DS2_REQUIRE(nnz >= 0 && B > 0, DS2_ERR_BAD_SHAPE, "ds2_dummy_sparse_labels: bad shape nnz=%d B=%d", nnz, B);
  if (nnz == 0) return DS2_OK;
  DS2_REQUIRE(nz_pos && nz_val && indices && values, DS2_ERR_BAD_ARG, "ds2_dummy_sparse_labels: null pointer");
  ds2::dummy_sparse_kernel<<<(B * nnz + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(nz_pos, nz_val, nnz, B, indices, values);
  DS2_LAUNCH_CHECK("dummy_sparse_kernel");
  ds2::count_launch();
  return DS2_OK;
}
