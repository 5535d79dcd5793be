// CTC loss + gradient w.r.t. unnormalised logits, fused with the softmax.
// Replaces tf.nn.ctc_loss as called at src/deepSpeech_NCHW.py:225 / src/deepSpeech.py:227
// (blank = K-1, time-major fp32 logits, labels already collapsed by the host shim).
//
// Kernel 1 (ctc_chain_kernel): grid (B, 2).  blockIdx.y==0 runs the alpha recursion forward in
// time, blockIdx.y==1 runs the beta recursion backward in time, concurrently on different SMs,
// one thread per lattice state, log-space fp32, re-normalised every RENORM steps so stored
// magnitudes stay small (offsets kept in fp64).  Tables go to a caller-provided scratch.
// Kernel 2 (ctc_grad_kernel): fully parallel over (t, b): one warp per logits row computes the
// softmax and subtracts the per-class occupancy sum_s exp(alpha+beta-logp).
#include <stdlib.h>

#include "common.cuh"

namespace ds2 {

static constexpr float NEGF = -1e30f;
static constexpr int RENORM = 16;

struct CtcWs {
  float* alpha;     // [B][T][Sp]
  float* beta;      // [B][T][Sp]
  double* coff;     // [B][2][NB]   cumulative renorm offsets (alpha, beta) per time block
  double* logp;     // [B]
  int* perm;        // [B][Lmax]    label positions sorted by class
  int* cls_off;     // [B][K+1]
  int Sp, NB;
};

__host__ __device__ inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

static size_t ctc_ws_layout(int T, int B, int K, int Lmax, char* base, CtcWs* ws) {
  int Sp = (2 * Lmax + 1 + 3) & ~3;
  int NB = (T + RENORM - 1) / RENORM + 1;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += align256(bytes); return o; };
  size_t o_alpha = take(sizeof(float) * (size_t)B * T * Sp);
  size_t o_beta = take(sizeof(float) * (size_t)B * T * Sp);
  size_t o_coff = take(sizeof(double) * (size_t)B * 2 * NB);
  size_t o_logp = take(sizeof(double) * (size_t)B);
  size_t o_perm = take(sizeof(int) * (size_t)B * (Lmax > 0 ? Lmax : 1));
  size_t o_cls = take(sizeof(int) * (size_t)B * (K + 1));
  if (ws) {
    ws->alpha = reinterpret_cast<float*>(base + o_alpha);
    ws->beta = reinterpret_cast<float*>(base + o_beta);
    ws->coff = reinterpret_cast<double*>(base + o_coff);
    ws->logp = reinterpret_cast<double*>(base + o_logp);
    ws->perm = reinterpret_cast<int*>(base + o_perm);
    ws->cls_off = reinterpret_cast<int*>(base + o_cls);
    ws->Sp = Sp;
    ws->NB = NB;
  }
  return off;
}

// log2(2^a + 2^b + 2^c): the recursion runs in base-2 log units (one MUFU.EX2 / MUFU.LG2 per term, no
// conversion multiplies on the dependency chain)
__device__ __forceinline__ float lse3(float a, float b, float c) {
  const float m = fmaxf(a, fmaxf(b, c));
  return m + __log2f(exp2f(a - m) + exp2f(b - m) + exp2f(c - m));
}
static constexpr float LOG2E = 1.4426950408889634f;
static constexpr double LN2 = 0.6931471805599453;

static constexpr int CTC_NSMAX = 4;   // lattice states per thread (S <= 4096)

}  // namespace ds2

#include "ds2_ctc_chain.cuh"
#include "ds2_ctc2.cuh"

namespace ds2 {

// NS = lattice states per thread (1 whenever 2*Lmax+1 <= 1024): the recursion is instruction-latency bound
// (ncu: ~70 % of issue slots lost to fixed-latency / MUFU dependencies), so the common case must not carry
// the predicated code of the multi-state case.
template <int NS>
__global__ void __launch_bounds__(1024, 1)
ctc_chain_kernel(const float* __restrict__ logits, int T, int B, int K, const int* __restrict__ labels, int Lmax,
                 const int* __restrict__ label_lens, const int* __restrict__ seq_lens, float* __restrict__ loss, CtcWs ws) {
  extern __shared__ __align__(16) float ctc_smem[];
  if (blockIdx.y == 0)
    ctc_chain<0, NS>(logits, T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws, ctc_smem);
  else
    ctc_chain<1, NS>(logits, T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws, ctc_smem);
}

// One warp per (t,b) row.  dynamic smem per warp: float e[Sp]
__global__ void __launch_bounds__(256)
ctc_grad_kernel(const float* __restrict__ logits, int T, int B, int K, const int* __restrict__ labels, int Lmax,
                const int* __restrict__ label_lens, const int* __restrict__ seq_lens, const float* __restrict__ dloss,
                float* __restrict__ grad, CtcWs ws, int rows_per_warp) {
  extern __shared__ __align__(16) float gsm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int b = blockIdx.y;
  float* e = gsm + (size_t)warp * ws.Sp;
  const int Tb = min(seq_lens[b], T);
  const int L = label_lens[b];
  const int S = 2 * L + 1;
  const int blank = K - 1;
  const double logp = ws.logp[b];
  const float up = dloss ? dloss[b] : 1.f;
  const int* perm = ws.perm + (size_t)b * Lmax;
  const int* cls = ws.cls_off + (size_t)b * (K + 1);
  const int t0 = (blockIdx.x * nwarps + warp) * rows_per_warp;
  for (int t = t0; t < min(t0 + rows_per_warp, T); ++t) {
    float* g = grad + ((size_t)t * B + b) * K;
    if (t >= Tb) {
      for (int k = lane; k < K; k += 32) g[k] = 0.f;
      continue;
    }
    // alpha rows written at step=t carry offset coff_a[t/RENORM]; beta rows written at
    // step=Tb-1-t carry coff_b[(Tb-1-t)/RENORM].
    const double ca = ws.coff[((size_t)b * 2 + 0) * ws.NB + t / RENORM];
    const double cb = ws.coff[((size_t)b * 2 + 1) * ws.NB + (Tb - 1 - t) / RENORM];
    const float c = (float)(ca + cb - logp);
    const float* ar = ws.alpha + ((size_t)b * T + t) * ws.Sp;
    const float* br = ws.beta + ((size_t)b * T + t) * ws.Sp;
    float blank_sum = 0.f;
    for (int s = lane; s < S; s += 32) {
      const float v = exp2f(fminf(ar[s] + br[s] + c, 1e-3f));     // tables, offsets and log p are in log2 units
      e[s] = v;
      if (!(s & 1)) blank_sum += v;
    }
    blank_sum = warp_sum(blank_sum);
    __syncwarp();
    const float* row = logits + ((size_t)t * B + b) * K;
    float m = -INFINITY;
    for (int k = lane; k < K; k += 32) m = fmaxf(m, row[k]);
    m = warp_max(m);
    float den = 0.f;
    for (int k = lane; k < K; k += 32) den += __expf(row[k] - m);
    den = warp_sum(den);
    const float inv = 1.f / den;
    for (int k = lane; k < K; k += 32) {
      float occ;
      if (k == blank) {
        occ = blank_sum;
      } else {
        occ = 0.f;
        for (int i = cls[k]; i < cls[k + 1]; ++i) occ += e[2 * perm[i] + 1];
      }
      g[k] = (__expf(row[k] - m) * inv - occ) * up;
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(256)
ctc_scale_grad_kernel(const float* __restrict__ grad, const float* __restrict__ dloss, float* __restrict__ out, long n, int B, int K) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    out[i] = grad[i] * dloss[(i / K) % B];
}

}  // namespace ds2

using namespace ds2;

extern "C" int ds2_ctc_scale_grad(const float* grad, const float* dloss, float* out, int T, int B, int K, void* stream) {
  DS2_REQUIRE(grad && dloss && out && T > 0 && B > 0 && K > 0, DS2_ERR_BAD_ARG, "ds2_ctc_scale_grad: bad argument");
  const long n = (long)T * B * K;
  long g = (n + 255) / 256;
  if (g > (long)num_sms() * 16) g = (long)num_sms() * 16;
  ctc_scale_grad_kernel<<<(int)g, 256, 0, static_cast<cudaStream_t>(stream)>>>(grad, dloss, out, n, B, K);
  DS2_LAUNCH_CHECK("ctc_scale_grad_kernel");
  count_launch();
  return DS2_OK;
}

namespace ds2 {
// the round-2 kernels take K <= 32 classes and up to 1024 lattice states (labels up to 511); anything else, and
// DS2_CTC_V1=1, runs the round-1 kernels.  One workspace size serves both.
static bool ctc2_takes(int K, int Lmax) { return K <= c2::CTC2_KP && 2 * Lmax + 1 <= 1024; }
static size_t ctc_ws_bytes_both(int T, int B, int K, int Lmax) {
  const size_t a = ctc_ws_layout(T, B, K, Lmax, nullptr, nullptr);
  const size_t b = ctc2_takes(K, Lmax) ? c2::ctc2_ws_layout(T, B, K, Lmax, nullptr, nullptr) : 0;
  return a > b ? a : b;
}

template <int NS, int W>
static int ctc2_launch_chain(int T, int B, int K, const int* labels, int Lmax, const int* label_lens, const int* seq_lens, float* loss,
                             const c2::Ws2& ws, size_t smem, cudaStream_t st) {
  static unsigned long long attr_done_devs = 0;
  if (first_use_on_device(&attr_done_devs))
    DS2_CUDA(cudaFuncSetAttribute(c2::ctc2_chain_kernel<NS, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
  c2::ctc2_chain_kernel<NS, W><<<dim3(B, 2), 32 * W, smem, st>>>(T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws);
  DS2_LAUNCH_CHECK("ctc2_chain_kernel");
  count_launch();
  return DS2_OK;
}

static int ctc2_launch(const float* logits, int T, int B, int K, const int* labels, int Lmax, const int* label_lens,
                       const int* seq_lens, const float* dloss, float* loss, float* grad, void* workspace, cudaStream_t st) {
  c2::Ws2 ws;
  c2::ctc2_ws_layout(T, B, K, Lmax, reinterpret_cast<char*>(workspace), &ws);
  const long rows = (long)T * B;
  c2::ctc2_rows_kernel<<<(unsigned)(((rows + c2::C2_ROWS - 1) / c2::C2_ROWS * 32 + 255) / 256), 256, 0, st>>>(logits, rows, K, ws.ly);
  DS2_LAUNCH_CHECK("ctc2_rows_kernel");
  count_launch();
  const size_t smem = sizeof(float) * (2 * c2::CTC2_RB * c2::CTC2_KP + 2 * (size_t)c2::CTC2_RB * ws.pitch + 16) + sizeof(int) * (size_t)(Lmax + 4);
  int rc = DS2_ERR_UNSUPPORTED;
#define DS2_C2_CASE(N, WW) if (ws.NS == N && ws.W == WW) rc = ctc2_launch_chain<N, WW>(T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws, smem, st);
  DS2_C2_CASE(2, 1) DS2_C2_CASE(2, 2) DS2_C2_CASE(2, 4) DS2_C2_CASE(4, 4) DS2_C2_CASE(6, 4) DS2_C2_CASE(8, 4)
#undef DS2_C2_CASE
  if (rc == DS2_ERR_UNSUPPORTED) set_error("ds2_ctc: no kernel for %d states per lane x %d warps", ws.NS, ws.W);
  if (rc) return rc;
  if (grad) {
    const int warps = 8;
    const size_t gsmem = sizeof(float) * ((size_t)warps * ws.pitch + ((Lmax + 3) & ~3) + ((K + 1 + 3) & ~3));
    const int rows_per_block = warps * c2::C2_GROWS;
    c2::ctc2_grad_kernel<<<dim3((T + rows_per_block - 1) / rows_per_block, B), warps * 32, gsmem, st>>>(T, B, K, Lmax, label_lens, seq_lens, dloss, grad, ws);
    DS2_LAUNCH_CHECK("ctc2_grad_kernel");
    count_launch();
  }
  return DS2_OK;
}
}  // namespace ds2

extern "C" size_t ds2_ctc_workspace_bytes(int T, int B, int K, int Lmax) {
  if (T <= 0 || B <= 0 || K <= 1 || Lmax < 0) return 0;
  return ds2::ctc_ws_bytes_both(T, B, K, Lmax);
}

namespace ds2 {
int ctc_launch(const float* logits, int T, int B, int K, const int* labels, int Lmax, const int* label_lens,
               const int* seq_lens, const float* dloss, float* loss, float* grad, void* workspace, size_t workspace_bytes,
               cudaStream_t st) {
  DS2_REQUIRE(logits && labels && label_lens && seq_lens && loss, DS2_ERR_BAD_ARG, "ds2_ctc: null pointer argument");
  DS2_REQUIRE(T > 0 && B > 0 && K > 1 && Lmax >= 0, DS2_ERR_BAD_SHAPE, "ds2_ctc: bad shape T=%d B=%d K=%d Lmax=%d", T, B, K, Lmax);
  CtcWs ws;
  ctc_ws_layout(T, B, K, Lmax, reinterpret_cast<char*>(workspace), &ws);
  const size_t need = ctc_ws_bytes_both(T, B, K, Lmax);
  DS2_REQUIRE(workspace && workspace_bytes >= need, DS2_ERR_WORKSPACE, "ds2_ctc: workspace too small (%zu < %zu bytes)",
              workspace_bytes, need);
  DS2_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, DS2_ERR_BAD_ARG, "ds2_ctc: workspace must be 256-byte aligned");
  if (!env().ctc_v1 && ctc2_takes(K, Lmax))
    return ctc2_launch(logits, T, B, K, labels, Lmax, label_lens, seq_lens, dloss, loss, grad, workspace, st);
  const int Smax = 2 * Lmax + 1;
  // states per thread: 1 measured fastest (one thread per lattice state keeps the per-step chain shortest; 2 and 4
  // halve / quarter the warps at the barrier but lengthen the serial work of a step); DS2_CTC_NS overrides for tuning
  int ns_pref = 1;
  { const int v = env().ctc_ns; if (v == 1 || v == 2 || v == 4) ns_pref = v; }
  int threads = (((Smax + ns_pref - 1) / ns_pref + 31) / 32) * 32;
  if (threads < 64) threads = 64;
  if (threads > 1024) threads = 1024;
  const size_t smem = ctc_smem_layout(T, Lmax, K, ws.Sp).bytes;
  DS2_REQUIRE(Smax <= 1024 * CTC_NSMAX, DS2_ERR_BAD_SHAPE, "ds2_ctc: Lmax=%d exceeds the supported label length %d", Lmax, (1024 * CTC_NSMAX - 1) / 2);
  DS2_REQUIRE(smem <= 200 * 1024, DS2_ERR_BAD_SHAPE, "ds2_ctc: T=%d / Lmax=%d need %zu bytes of shared memory (> 200 KB)", T, Lmax, smem);
  static unsigned long long attr_done_devs = 0;
  if (first_use_on_device(&attr_done_devs)) {
    DS2_CUDA(cudaFuncSetAttribute(ctc_chain_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    DS2_CUDA(cudaFuncSetAttribute(ctc_chain_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    DS2_CUDA(cudaFuncSetAttribute(ctc_chain_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    DS2_CUDA(cudaFuncSetAttribute(ctc_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  }
  const int ns = (Smax + threads - 1) / threads;
  if (ns <= 1) ctc_chain_kernel<1><<<dim3(B, 2), threads, smem, st>>>(logits, T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws);
  else if (ns == 2) ctc_chain_kernel<2><<<dim3(B, 2), threads, smem, st>>>(logits, T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws);
  else ctc_chain_kernel<4><<<dim3(B, 2), threads, smem, st>>>(logits, T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws);
  DS2_LAUNCH_CHECK("ctc_chain_kernel");
  count_launch();
  if (grad) {
    const int warps = 8;
    const int rows_per_warp = 1;
    size_t gsmem = sizeof(float) * (size_t)warps * ws.Sp;
    DS2_REQUIRE(gsmem <= 64 * 1024, DS2_ERR_BAD_SHAPE, "ds2_ctc: Lmax=%d too large for the gradient kernel", Lmax);
    dim3 grid((T + warps * rows_per_warp - 1) / (warps * rows_per_warp), B);
    ctc_grad_kernel<<<grid, warps * 32, gsmem, st>>>(logits, T, B, K, labels, Lmax, label_lens, seq_lens, dloss, grad, ws,
                                                     rows_per_warp);
    DS2_LAUNCH_CHECK("ctc_grad_kernel");
    count_launch();
  }
  return DS2_OK;
}
}  // namespace ds2

extern "C" int ds2_ctc_loss_fwd_bwd(const float* logits, int T, int B, int K, const int* labels, int Lmax,
                                    const int* label_lens, const int* seq_lens, const float* dloss, float* loss, float* grad,
                                    void* workspace, size_t workspace_bytes, void* stream) {
  return ctc_launch(logits, T, B, K, labels, Lmax, label_lens, seq_lens, dloss, loss, grad, workspace, workspace_bytes,
                    static_cast<cudaStream_t>(stream));
}

// ------------------------------------------------------------------ allocation-free variant
// Some callers cannot give this library a scratch buffer because they cannot allocate one at the point of the
// call: a host framework that has handed the whole of HBM to its own arena (cudaMalloc then has nothing left
// to return), or a latency-critical step loop in which cudaMalloc/cudaFree would stall the device.  For them
// the alpha/beta tables live in a fixed array that is part of the loaded module: every GPU that loads the
// library carries these DS2_CTC_NOALLOC_MAX_BYTES until the process exits, used or not.
#include <mutex>

#define DS2_CTC_NOALLOC_MAX_BYTES (64ull << 20)
static __device__ __align__(256) unsigned char g_ctc_noalloc_scratch[DS2_CTC_NOALLOC_MAX_BYTES];

// The array's name is only a handle on the host side; the address kernels can use is fetched once per device,
// with that device current, and remembered.
static int ctc_noalloc_scratch(void** out) {
  static std::mutex mu;
  static void* dev_ptr[64] = {};
  int dev = -1;
  DS2_CUDA(cudaGetDevice(&dev));
  DS2_REQUIRE(dev >= 0 && dev < 64, DS2_ERR_CUDA, "ds2_ctc_loss_fwd_bwd_noalloc: unexpected device ordinal %d", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!dev_ptr[dev]) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_ctc_noalloc_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_ctc_noalloc_scratch)");
    DS2_REQUIRE(ptr != nullptr, DS2_ERR_CUDA, "ds2_ctc_loss_fwd_bwd_noalloc: no device address for the built-in scratch");
    dev_ptr[dev] = ptr;
  }
  *out = dev_ptr[dev];
  return DS2_OK;
}

// Also the set-up call: it looks up the array's device address on the CURRENT device (no allocation; with lazy module
// loading this is the moment the driver backs the array with memory).  A caller whose own allocator is about to claim
// the rest of HBM, or who is about to begin a stream capture, calls this first, once per device.
extern "C" size_t ds2_ctc_noalloc_max_bytes(void) {
  void* scratch = nullptr;
  (void)ctc_noalloc_scratch(&scratch);      // a failure is reported again, with its message, by the first compute call
  return (size_t)DS2_CTC_NOALLOC_MAX_BYTES;
}

// Same computation and same kernels as ds2_ctc_loss_fwd_bwd; only the home of the scratch differs.
//  * The first call on each device looks up the scratch address (no allocation, but a runtime query): make that
//    first call before any stream capture is begun on that device.
//  * Problems whose tables do not fit are refused (DS2_ERR_WORKSPACE, message names both sizes); nothing is
//    launched in that case.  ds2_ctc_workspace_bytes(T,B,K,Lmax) <= ds2_ctc_noalloc_max_bytes() is the test.
//  * All calls on a device share the one array and it keeps whatever the previous call left there.  Nothing in
//    it has to start at zero - every cell a call reads it has written first - so no clearing is done; but two
//    calls whose kernels overlap on the GPU (different streams, no dependency) overwrite each other's tables.
extern "C" int ds2_ctc_loss_fwd_bwd_noalloc(const float* logits, int T, int B, int K, const int* labels, int Lmax,
                                            const int* label_lens, const int* seq_lens, const float* dloss, float* loss,
                                            float* grad, void* stream) {
  DS2_REQUIRE(T > 0 && B > 0 && K > 1 && Lmax >= 0, DS2_ERR_BAD_SHAPE, "ds2_ctc_loss_fwd_bwd_noalloc: bad shape T=%d B=%d K=%d Lmax=%d",
              T, B, K, Lmax);
  const size_t need = ctc_ws_bytes_both(T, B, K, Lmax);
  DS2_REQUIRE(need <= (size_t)DS2_CTC_NOALLOC_MAX_BYTES, DS2_ERR_WORKSPACE,
              "ds2_ctc_loss_fwd_bwd_noalloc: T=%d B=%d Lmax=%d needs %zu bytes of scratch, the built-in array holds %zu; "
              "use ds2_ctc_loss_fwd_bwd with a caller-provided workspace",
              T, B, Lmax, need, (size_t)DS2_CTC_NOALLOC_MAX_BYTES);
  void* scratch = nullptr;
  int rc = ctc_noalloc_scratch(&scratch);
  if (rc) return rc;
  return ctc_launch(logits, T, B, K, labels, Lmax, label_lens, seq_lens, dloss, loss, grad, scratch,
                    (size_t)DS2_CTC_NOALLOC_MAX_BYTES, static_cast<cudaStream_t>(stream));
}
