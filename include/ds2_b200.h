/* ds2_b200 - C ABI of the B200-native DeepSpeech2 training hot path.
 *
 * The reference (yxlao/deepSpeech) has no FFI: its hot path is a set of Python call sites into
 * TensorFlow's C++ CPU kernels.  Each entry point below replaces the TF op(s) reached from the
 * cited reference line(s); a reference maintainer would bind them with ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void*; work is stream-ordered, nothing synchronises;
 *   - the library borrows the caller's buffers for the duration of the call and allocates nothing
 *     itself: scratch is passed in explicitly (ds2_*_workspace_bytes tells how much);
 *   - return value: DS2_OK or a DS2_ERR_* code; ds2_last_error() returns the message of the last
 *     failure on the calling thread;
 *   - dtypes: DS2_F32 = 0, DS2_BF16 = 1;
 *   - activations are channel-last: conv tensors [B,T,F,C], recurrent tensors time-major [T,B,*].
 */
#ifndef DS2_B200_H_
#define DS2_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { DS2_F32 = 0, DS2_BF16 = 1 };
enum {
  DS2_OK = 0,
  DS2_ERR_BAD_ARG = 1,      /* null pointer, misaligned pointer, bad enum                         */
  DS2_ERR_BAD_SHAPE = 2,    /* negative / inconsistent / unsupported dimensions                   */
  DS2_ERR_WORKSPACE = 3,    /* scratch buffer missing or too small                                */
  DS2_ERR_UNSUPPORTED = 4,  /* shape/dtype combination this build has no kernel for               */
  DS2_ERR_CUDA = 5          /* a CUDA runtime/driver call failed (message has the CUDA error)     */
};

const char* ds2_last_error(void);
int ds2_version(void);
/* Re-reads the DS2_* debugging switches from the environment (they are otherwise read once per process).  No
 * counterpart in the reference; used by the tests that flip a switch inside one process. */
void ds2_debug_reload_env(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
long long ds2_launch_count(void);
/* Caps the number of CTAs that grid-sized (persistent / grid-stride) kernels launched afterwards from THIS host
 * thread use: n SMs' worth, 0 = the whole device.  Returns the previous value.  No counterpart in the reference:
 * the host layer uses it to run weight-gradient GEMMs on a low-priority side stream next to a recurrence kernel
 * that occupies a fixed set of SMs (deepspeech_b200/ops.py, recurrent stack backward). */
int ds2_set_sm_limit(int n);

/* ---- CTC ------------------------------------------------------------------------------------
 * tf.nn.ctc_loss(labels, inputs=float32 logits [T,B,K], sequence_length, time_major=True) as
 * called at src/deepSpeech_NCHW.py:225 and src/deepSpeech.py:227: blank = K-1, softmax inside,
 * loss[b] = -log p(label_b | logits_b[:seq_lens[b]]), grad = d(sum_b dloss[b]*loss[b])/d logits
 * (zero for t >= seq_lens[b]).  `labels` is [B,Lmax] int32, row b holds label_lens[b] class ids
 * in [0,K-2] (already collapsed/terminated by the host shim, src/deepSpeech_dummy.py:89-98 +
 * TF's preprocess_collapse_repeated).  dloss may be NULL (= all ones); grad may be NULL.
 * The caller must have checked seq_lens[b] >= label_lens[b] + repeats (TF raises otherwise). */
size_t ds2_ctc_workspace_bytes(int T, int B, int K, int Lmax);
int ds2_ctc_loss_fwd_bwd(const float* logits, int T, int B, int K, const int* labels, int Lmax, const int* label_lens,
                         const int* seq_lens, const float* dloss, float* loss, float* grad, void* workspace,
                         size_t workspace_bytes, void* stream);
/* Allocation-free form of ds2_ctc_loss_fwd_bwd.  It has NO counterpart in the reference: it exists for
 * callers of this library that cannot obtain scratch memory at the point of the call - a host framework that
 * has handed all of HBM to its own arena so that cudaMalloc has nothing left to give, or a step loop that
 * cannot afford the device stall of cudaMalloc/cudaFree.  Same arguments minus the workspace; the alpha/beta
 * tables live in a fixed array inside the library (ds2_ctc_noalloc_max_bytes() = 64 MiB, resident on every GPU
 * that loads the library until the process exits).  Limits and rules:
 *   - refused with DS2_ERR_WORKSPACE when ds2_ctc_workspace_bytes(T,B,K,Lmax) exceeds that maximum
 *     (the reference's largest dummy bucket at B=32 - T2=366, 209 labels - needs 39.4 MB and fits);
 *   - ds2_ctc_noalloc_max_bytes() is also the set-up call: on the CURRENT device it resolves the array's device
 *     address (a runtime query, no allocation; with lazy module loading it is also the moment the driver backs
 *     the array with memory).  Call it once per device BEFORE your own allocator claims the rest of HBM and
 *     before any stream capture; if you do not, the first compute call on the device does the same set-up;
 *   - the entry point itself calls no cudaMalloc/cudaFree or any allocating routine;
 *   - all calls share the array and it keeps the previous call's contents; each call overwrites every cell
 *     it later reads, so nothing is cleared, but calls that overlap on the GPU (two streams) race on it. */
size_t ds2_ctc_noalloc_max_bytes(void);
int ds2_ctc_loss_fwd_bwd_noalloc(const float* logits, int T, int B, int K, const int* labels, int Lmax,
                                 const int* label_lens, const int* seq_lens, const float* dloss, float* loss, float* grad,
                                 void* stream);
/* out[t,b,k] = grad[t,b,k] * dloss[b]: applies the upstream d/dloss_b (e.g. 1/B of reduce_mean,
 * src/deepSpeech_NCHW.py:226) to a gradient computed with dloss = NULL. */
int ds2_ctc_scale_grad(const float* grad, const float* dloss, float* out, int T, int B, int K, void* stream);

/* ---- GEMM -----------------------------------------------------------------------------------
 * tf.matmul call sites: src/custom_ops.py:61 (W.x, batched over all timesteps), :67 (U.h, see
 * ds2_rnn_*), :249-251 (_linear), src/deepSpeech_NCHW.py:196 (FC), and their gradients
 * (src/deepSpeech_train.py:287).
 * ds2_gemm_bf16: tcgen05/TMA kernel.  C[M,N] = alpha*A*B^T + bias[n] + beta*C.  A is [M,K]
 *   row-major (a_mn_major=0, lda = row pitch) or [K,M] row-major (a_mn_major=1); same for B with N.
 *   Pitches in elements, multiples of 8; bases 16-byte aligned.  c_dtype DS2_F32 or DS2_BF16.
 * ds2_gemm_generic: any strides / dtypes, CUDA cores, fp32 accumulation (small or odd shapes).
 *   C[m,n] = alpha * sum_k A[m*sam + k*sak] * B[n*sbn + k*sbk] + bias[n] + beta*C[m*ldc+n]. */
int ds2_gemm_bf16(const void* A, long lda, int a_mn_major, const void* B, long ldb, int b_mn_major, void* C, int c_dtype,
                  long ldc, int M, int N, int K, float alpha, float beta, const float* bias, void* stream);
/* split-K variant for short-and-wide products (weight gradients: K = T*B or B*To*Fo rows): C is zeroed,
 * then every K-split red.add's its partial tile.  fp32 C, beta = 0.  splits <= 0: chosen to fill the SMs. */
int ds2_gemm_bf16_splitk(const void* A, long lda, int a_mn_major, const void* B, long ldb, int b_mn_major, float* C, long ldc,
                         int M, int N, int K, float alpha, const float* bias, int splits, void* stream);
int ds2_gemm_generic(const void* A, int a_dtype, long sam, long sak, const void* B, int b_dtype, long sbn, long sbk, void* C,
                     int c_dtype, long ldc, int M, int N, int K, float alpha, float beta, const float* bias, void* stream);

/* ---- recurrent layers -----------------------------------------------------------------------
 * CustomRNNCell2.__call__ (src/custom_ops.py:48-72) unrolled over time by
 * tf.nn.bidirectional_dynamic_rnn (src/custom_ops.py:91-93, src/deepSpeech.py:174-183):
 *   h[d,t] = t < lens[b] ? min(max(scale*wx[t,b,d*H:] + U[d] h[d,t-/+1] + bias[d], 0), cap) : 0
 * wx   [T,B,ND*H] f32  - input projection W.x for all timesteps (scale = seq_batch_norm's
 *                        1/sqrt(1+1e-5), src/custom_ops.py:184-199 as written);
 * u    [ND,H,H], bias [ND,H] f32, lens [B] int32, hs [ND,T,B,H] out; direction d runs backward in
 * time when bit d of reverse_mask is set.  dtype selects the path: DS2_F32 = per-step CUDA-core
 * kernels, u/hs fp32; DS2_BF16 = persistent tcgen05 kernel, u/hs bf16 (see
 * ds2_rnn_relu_persistent_supported).
 * Backward: dout [T,B,H] f32 is the gradient w.r.t. sum_d hs[d]; dz [T,B,ND*H] (f32|bf16) receives
 * the gradient w.r.t. the pre-activation, from which the caller forms dW, dU, dx by GEMM; dbias [ND,H] f32 (may be
 * NULL) receives the bias gradient sum_{t,b} dz, accumulated inside the persistent kernel. */
int ds2_rnn_relu_fwd(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs, int dtype,
                     int T, int B, int H, int ND, int reverse_mask, float cap, void* workspace, size_t workspace_bytes,
                     void* stream);
/* ds2_rnn_relu_fwd on the bf16 tensor-core path with one more output: hs_f32 [ND,T,B,H] (may be NULL) receives the
 * hidden states BEFORE their rounding to bf16, so that the caller can form the layer output fw + bw
 * (src/custom_ops.py:95) from unrounded values - one rounding per layer instead of three. */
int ds2_rnn_relu_fwd_ex(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs, float* hs_f32,
                        int T, int B, int H, int ND, int reverse_mask, float cap, void* workspace, size_t workspace_bytes,
                        void* stream);
int ds2_rnn_relu_bwd(const float* dout, const void* u, const int* lens, const void* hs, void* dz, float* dbias, int dtype, int T,
                     int B, int H, int ND, int reverse_mask, float cap, void* workspace, size_t workspace_bytes, void* stream);
/* ds2_rnn_relu_bwd on the bf16 path with the transposed recurrent weights supplied by the caller: ut [ND,H,H] bf16,
 * ut[d][k][j] = u[d][j][k] (the operand the backward contraction U^T dz keeps resident).  Weights change once per
 * optimizer step, so a caller can transpose them once instead of once per layer and step; with ut == NULL the call
 * transposes u into the workspace like ds2_rnn_relu_bwd (u may be NULL when ut is given). */
int ds2_rnn_relu_bwd_ex(const float* dout, const void* u, const void* ut, const int* lens, const void* hs, void* dz, float* dbias, int T,
                        int B, int H, int ND, int reverse_mask, float cap, void* workspace, size_t workspace_bytes, void* stream);
size_t ds2_rnn_workspace_bytes(int T, int B, int H, int ND, int dtype);
int ds2_rnn_relu_persistent_supported(int B, int H, int ND);

/* GRU (gate order r,z,n; cuDNN / torch.nn.GRU form).  NOT in the reference: BASELINE
 * configs 3 and 5 ask for it, the build defines it.  wx [T,B,ND*3H], u [ND,3H,H], bw/bu [ND,3H],
 * hs [ND,T,B,H], gates [ND,T,B,4H] (r,z,n,U_n h+bu_n saved for backward), dgi/dgh [T,B,ND*3H]. */
int ds2_gru_fwd(const float* wx, float scale, const void* u, const float* bw, const float* bu, const int* lens, void* hs,
                void* gates, int dtype, int T, int B, int H, int ND, int reverse_mask, void* stream);
int ds2_gru_bwd(const float* dout, const void* u, const int* lens, const void* hs, const void* gates, void* dgi, void* dgh,
                float* dhz_scratch, int dtype, int T, int B, int H, int ND, int reverse_mask, void* stream);
/* Persistent tcgen05 form of the same GRU (one launch per layer and pass, U resident in TMEM): u, hs, gates, dgi, dgh
 * are bf16, wx/dout/bw/bu fp32, same layouts as above.  ds2_gru_persistent_supported says whether (B,H,ND) fits
 * (B = 16 or a multiple of 32, H%8==0; H <= 1024 runs both directions in one launch, wider layers use an 8-CTA K split
 * and one launch per direction);
 * workspace = ds2_gru_workspace_bytes (step counters + U^T for the backward), the caller owns it.  dbw / dbu [ND,3H] f32
 * (both or neither, may be NULL) receive the bias gradients sum_{t,b} dgi and sum_{t,b} dgh. */
int ds2_gru_persistent_supported(int B, int H, int ND);
size_t ds2_gru_workspace_bytes(int T, int B, int H, int ND);
int ds2_gru_persistent_fwd(const float* wx, float scale, const void* u, const float* bw, const float* bu, const int* lens,
                           void* hs, void* gates, int T, int B, int H, int ND, int reverse_mask, void* workspace,
                           size_t workspace_bytes, void* stream);
int ds2_gru_persistent_bwd(const float* dout, const void* u, const int* lens, const void* hs, const void* gates, void* dgi,
                           void* dgh, float* dbw, float* dbu, int T, int B, int H, int ND, int reverse_mask, void* workspace,
                           size_t workspace_bytes, void* stream);

/* ---- conv front-end -------------------------------------------------------------------------
 * tf.nn.conv2d + bias_add at src/deepSpeech_NCHW.py:120-127,:142-151 (src/deepSpeech.py:120-126,
 * :143-149): VALID cross-correlation, HWIO filters w[KT,KF,Ci,Co], strides (st,sf).
 * x [B,Ti,Fi,Ci] (f32|bf16), y [B,To,Fo,Co] f32, To=(Ti-KT)/st+1, Fo=(Fi-KF)/sf+1. */
int ds2_conv2d_fwd(const void* x, int x_dtype, const float* w, const float* bias, float* y, int B, int Ti, int Fi, int Ci,
                   int Co, int KT, int KF, int st, int sf, void* stream);
int ds2_conv2d_dgrad(const float* dy, const float* w, float* dx, int B, int Ti, int Fi, int Ci, int Co, int KT, int KF,
                     int st, int sf, void* stream);
int ds2_conv2d_wgrad(const void* x, int x_dtype, const float* dy, float* dw, float* dbias, int B, int Ti, int Fi, int Ci,
                     int Co, int KT, int KF, int st, int sf, void* stream);

/* Tensor-core route for the same convolutions: patch matrix P[M = B*To*Fo, ldp] (bf16, k = (kt*KF+kf)*Ci+ci,
 * zero padded to ldp % 8 == 0) gathered from the fp32 channel-last input, optionally applying the previous
 * layer's batch-norm + clipped ReLU on the fly (mean == NULL: plain gather).  Then y = P.W via ds2_gemm_bf16,
 * dW = P^T.dy via ds2_gemm_bf16_splitk, dP = dy.W^T via ds2_gemm_bf16 and ds2_col2im sums dP back onto the
 * input grid (dgrad).  ds2_col2im needs Ci % 8 == 0. */
int ds2_im2col(const float* x, const float* mean, const float* rstd, const float* gamma, const float* beta, float cap,
               int clip, void* P, long ldp, int B, int Ti, int Fi, int Ci, int KT, int KF, int st, int sf, void* stream);
int ds2_col2im(const void* dP, long ldp, float* dx, int B, int Ti, int Fi, int Ci, int KT, int KF, int st, int sf,
               void* stream);

/* Fused implicit-GEMM convolution of a one-channel input into 32 feature maps (the reference's conv1: filter
 * [20,5,1,32], strides (2,2), src/deepSpeech_NCHW.py:110-119) - no patch matrix: the CTA builds the patch rows of a
 * 128-position tile in shared memory as the tcgen05 A operand (deepspeech_b200/csrc/ds2_conv1_fused.cu).
 *   x f32 [B,Ti,Fi], w f32 HWIO [KT,KF,1,32], bias f32 [32] or NULL, y f32 [B,To,Fo,32];
 *   sums (optional) f64 [64]: per-channel sum and sum of squares of y from the epilogue, for ds2_bn_finalize;
 *   wgrad: dy f32 [B,To,Fo,32] -> dw f32 HWIO (every element written), db f32 [32] (optional) = column sums of dy.
 * Operands are rounded to bf16, accumulation is fp32. */
int ds2_conv1_fused_supported(int Ci, int Co, int Ti, int Fi, int KT, int KF, int st, int sf);
int ds2_conv1_fused_fwd(const float* x, const float* w, const float* bias, float* y, double* sums, int B, int Ti, int Fi, int KT, int KF,
                        int st, int sf, void* stream);
int ds2_conv1_fused_wgrad(const float* x, const float* dy, float* dw, float* db, int B, int Ti, int Fi, int KT, int KF, int st, int sf,
                          void* stream);

/* Implicit-GEMM form of the same convolution for 32-channel layers with unit frequency stride (the reference's
 * conv2: filter [10,5,32,32], strides (2,1)): no patch matrix, the tensor cores read overlapping windows of the
 * channel-last activation through a strided TMA view (deepspeech_b200/csrc/ds2_conv_window.cu).
 *   a    bf16 [B,Ti,Fi,32]  - the layer input (for conv2: clip(BN(conv1)), written by ds2_bn_clip_fwd); the buffer
 *                             must extend 64 bytes past its last element (the last window reads them, unused);
 *   wt   bf16 [32][KT*KF*32] - filters as wt[co][(kt,kf,ci)],  wd bf16 [32][KT*KF*32] - wd[ci][(kt,KF-1-kf,co)],
 *                             both made from the HWIO fp32 filter by ds2_conv_window_weights;
 *   dyp  bf16 [B, Rdy, Fo+2*(KF-1), 32] - dL/dy with KF-1 zero positions on both sides of every row, `ptop` zero rows
 *                             above the To rows of every batch item and zero rows below them up to Rdy rows per item
 *                             (both from ds2_conv_window_dy_layout; written by ds2_conv_window_pad_cast): the kernels
 *                             pack several rows into one 128-row MMA tile and read the rows a filter tap reaches
 *                             outside [0,To) as zeros; 64 bytes of slack after the last element as for `a`;
 *   y    f32 [B,To,Fo,32], da f32 [B,Ti,Fi,32] (every element written), dw f32 HWIO [KT,KF,32,32];
 *   sums (optional) f64 [64]: per-channel sum and sum of squares of y from the forward epilogue, for ds2_bn_finalize.
 * ds2_conv_window_supported tells whether a shape is taken. */
int ds2_conv_window_supported(int Ci, int Co, int Fi, int KT, int KF, int sf);
int ds2_conv_window_weights(const float* w, void* wt, void* wd, int KT, int KF, void* stream);
int ds2_conv_window_dy_layout(int Ti, int Fi, int KT, int KF, int st, int* ptop, int* rows_per_item);
int ds2_conv_window_pad_cast(const float* dy, void* dyp, int B, int Ti, int Fi, int KT, int KF, int st, void* stream);
int ds2_conv_window_fwd(const void* a, const void* wt, const float* bias, float* y, double* sums, int B, int Ti, int Fi, int KT, int KF,
                        int st, void* stream);
int ds2_conv_window_dgrad(const void* dyp, const void* wd, float* da, int B, int Ti, int Fi, int KT, int KF, int st, void* stream);
int ds2_conv_window_wgrad(const void* a, const void* dyp, float* dw, int B, int Ti, int Fi, int KT, int KF, int st, void* stream);

/* batch_norm2 / batch_norm (src/custom_ops.py:107-160, :163-181; training mode, biased batch
 * variance over all rows of the channel-last [R,C] tensor) fused with relux(.., capping)
 * (src/custom_ops.py:99-104).  C <= 256.  sums_scratch: 2*C doubles.
 * The activation-side tensor (out / dout) is addressed as b*sb + t*st + f*sf + c*sc so it can be
 * written straight into the RNN layout [T,B,C*F] (src/deepSpeech_NCHW.py:166-168) or [T,B,F*C]
 * (src/deepSpeech.py:162-164). */
int ds2_bn_stats(const float* x, long R, int C, float eps, double* sums_scratch, float* mean, float* rstd, void* stream);
/* mean / rstd from per-channel sums a producer kernel accumulated in its epilogue (sums[c] = sum of x, sums[C + c] = sum
 * of x^2 over `count` values per channel): what ds2_bn_stats computes, without its pass over x. */
int ds2_bn_finalize(const double* sums, double count, int C, float eps, float* mean, float* rstd, void* stream);
int ds2_bn_clip_fwd(const float* x, int B, int T, int F, int C, const float* mean, const float* rstd, const float* gamma,
                    const float* beta, float cap, int clip, void* out, int out_dtype, long sb, long st, long sf, long sc,
                    void* stream);
int ds2_bn_clip_bwd(const float* x, int B, int T, int F, int C, const float* mean, const float* rstd, const float* gamma,
                    const float* beta, float cap, int clip, const void* dout, int dout_dtype, long sb, long st, long sf,
                    long sc, double* sums_scratch, float* dx, float* dgamma, float* dbeta, void* stream);

/* ---- element-wise helpers -------------------------------------------------------------------
 * relux: src/custom_ops.py:99-104.  add2: fw+bw sum, src/custom_ops.py:95, src/deepSpeech.py:185. */
int ds2_relux_fwd(const float* x, float* y, long n, float cap, int has_cap, void* stream);
int ds2_relux_bwd(const float* x, const float* dy, float* dx, long n, float cap, int has_cap, void* stream);
/* y = s*x: seq_batch_norm as written, src/custom_ops.py:184-199 (x / sqrt(1+1e-5)). */
int ds2_scale(const float* x, float s, float* y, long n, void* stream);
int ds2_add2(const void* a, const void* b, int in_dtype, void* out, int out_dtype, long n, void* stream);
int ds2_cast(const void* in, int in_dtype, void* out, int out_dtype, long n, void* stream);
int ds2_colsum(const void* x, int dtype, long R, int N, long ld, float* out, int accumulate, void* stream);
int ds2_transpose(const void* in, int in_dtype, long R, long C, long ldi, void* out, int out_dtype, long ldo, void* stream);

/* ---- optimizer step (the step right after the hot path) -----------------------
 * tf.train.AdamOptimizer(lr).apply_gradients + tf.train.ExponentialMovingAverage(decay, global_step).apply
 * (src/deepSpeech_train.py:430, :457-465) for ALL tensors in one launch.  `table` is a device array of n_tensors rows
 * of 7 int64: {w, g, m, v, shadow (0 = none) device pointers (fp32), numel, index of the tensor's first chunk}, chunks
 * of ds2_adam_ema_chunk() elements numbered consecutively over the tensors (n_chunks in total).  lr is the decayed
 * learning rate of this step (tf.train.exponential_decay, staircase, :231-250), step = 1 for the first update,
 * ema_decay = min(decay, (1+n)/(10+n)) as TF computes it; the shadow follows the UPDATED weight; g is scaled by
 * grad_scale first. */
int ds2_adam_ema_chunk(void);
int ds2_adam_ema_step(const void* table, int n_tensors, int n_chunks, float lr, float beta1, float beta2, float eps, long step,
                      float ema_decay, float grad_scale, void* stream);

/* ---- evaluation-side consumers of the logits ----------------------------------
 * ds2_ctc_greedy_decode: tf.nn.ctc_greedy_decoder(log_softmax(logits), seq_lens) as used at
 *   src/deepSpeech_test.py:209-213: per-frame argmax over K <= 32 classes (lowest index on ties), runs collapsed,
 *   blank = K-1 removed.  logits f32 [T,B,K]; decoded int32 [B,T] (row b holds decoded_lens[b] class ids).
 * ds2_edit_distance: Levenshtein distance of n pairs of int32 sequences (rows of a [n,lda] / b [n,ldb] with lengths
 *   a_lens / b_lens <= maxlen <= 4000): the numerator of the character error rate, src/deepSpeech_test.py:126-130. */
int ds2_ctc_greedy_decode(const float* logits, int T, int B, int K, const int* seq_lens, int* decoded, int* decoded_lens,
                          void* stream);
int ds2_edit_distance(const int* a, int lda, const int* a_lens, const int* b, int ldb, const int* b_lens, int n, int maxlen, int* out,
                      void* stream);

/* ---- label pre-processing on the device ------------------------------------------
 * What tf.nn.ctc_loss does with the SparseTensor labels of src/deepSpeech_train.py:437 before the lattice: per batch
 * row, in index order, collapse adjacent repeats when asked, stop at the first label >= num_classes-1.  indices [n,2]
 * int64 sorted row-major, values [n] int32 (both on the device); labels [B,Lcap] (zero padded), label_lens, repeats
 * (adjacent equal labels that remain: the lattice needs that many extra frames) and status [B] come back: bit 0 = a
 * valid label follows a terminator (TF raises), bit 1 = columns not sorted, bit 2 = more than Lcap labels.
 * The host form with the reference's exceptions is deepspeech_b200.ops.ctc_prepare_labels. */
int ds2_ctc_labels_from_sparse(const long long* indices, const int* values, int n, int B, int Lcap, int num_classes,
                               int collapse_repeated, int* labels, int* label_lens, int* repeats, int* status, void* stream);
/* The label half of the dummy batch producer on the device (src/deepSpeech_dummy.py:89-98, :115-118): one label row shared
 * by all B utterances, zeros dropped; nz_pos / nz_val [nnz] are the row's non-zero positions and values (device);
 * indices [B*nnz,2] int64 and values [B*nnz] int32 are the SparseTensor of src/deepSpeech_train.py:437. */
int ds2_dummy_sparse_labels(const int* nz_pos, const int* nz_val, int nnz, int B, long long* indices, int* values, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DS2_B200_H_ */
