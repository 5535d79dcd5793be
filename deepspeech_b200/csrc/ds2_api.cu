// Error plumbing, device queries and the recurrent-layer dispatch of the C ABI.
#include <stdarg.h>
#include <stdlib.h>

#include <atomic>

#include <mutex>

#include "common.cuh"

namespace ds2 {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return DS2_ERR_CUDA;
}
static thread_local int g_sm_limit = 0;     // ds2_set_sm_limit: 0 = whole device
int num_sms() {
  static int n[64] = {0};
  int dev = 0;
  int v = 148;
  if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64) {
    if (!n[dev]) {
      int q = 0;
      if (cudaDeviceGetAttribute(&q, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || q <= 0) q = 148;
      n[dev] = q;
    }
    v = n[dev];
  }
  return (g_sm_limit > 0 && g_sm_limit < v) ? g_sm_limit : v;
}
bool first_use_on_device(unsigned long long* flags) {
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;     // unknown device: set the attributes again
  std::lock_guard<std::mutex> lock(mu);
  const unsigned long long bit = 1ull << dev;
  if (*flags & bit) return false;
  *flags |= bit;
  return true;
}
int set_sm_limit(int n) {
  const int prev = g_sm_limit;
  g_sm_limit = n > 0 ? n : 0;
  return prev;
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static EnvSwitches g_env;
static std::once_flag g_env_once;
static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  if (!v) return dflt;
  if (!*v) return 1;
  return atoi(v);
}
static int env_flag(const char* name) { return getenv(name) ? 1 : 0; }
void reload_env() {
  EnvSwitches e = {};
  e.rnn_no_coop = env_flag("DS2_RNN_NO_COOP");
  e.rnn_exp = env_int("DS2_RNN_EXP", 0);
  e.rnn_fenced = env_flag("DS2_RNN_FENCED");
  e.rnn_slab_f32 = env_flag("DS2_RNN_SLAB_F32");
  e.rnn_prof = env_flag("DS2_RNN_PROF");
  e.rnn_v1 = env_flag("DS2_RNN_V1");
  {
    const char* m = getenv("DS2_RNN_POLL_MARGIN");
    e.rnn_poll_margin = m ? atoi(m) : -1;       // -1: the kernel's own table
  }
  e.gru_no_bulk = env_flag("DS2_GRU_NO_BULK");
  e.gemm_no_pair = env_flag("DS2_GEMM_NO_PAIR");
  e.ctc_ns = env_int("DS2_CTC_NS", 0);
  e.ctc_v1 = env_flag("DS2_CTC_V1");
  // Nsight Compute / Systems and compute-sanitizer inject a library into the target process and announce it in the
  // environment (observed under ncu 2025.2: NV_COMPUTE_PROFILER_PERFWORKS_DIR, NV_NSIGHT_INJECTION_*,
  // NV_TPS_LAUNCH_TOKEN; CUDA_INJECTION64_PATH is the documented generic hook)
  e.profiler = (getenv("NV_COMPUTE_PROFILER_PERFWORKS_DIR") || getenv("CUDA_INJECTION64_PATH") || getenv("NV_TPS_LAUNCH_TOKEN") ||
                getenv("NV_NSIGHT_INJECTION_PORT_BASE") || getenv("NSYS_PROFILING_SESSION_ID") || getenv("NV_SANITIZER_INJECTION_PORT_BASE"))
                   ? 1 : 0;
  g_env = e;
}
const EnvSwitches& env() {
  std::call_once(g_env_once, reload_env);
  return g_env;
}
bool use_cooperative_launch() { return !env().rnn_no_coop && !env().profiler; }

// implemented in ds2_rnn_generic.cu / ds2_rnn_persist.cu
int rnn_relu_fwd_generic(const float* wx, float scale, const float* u, const float* bias, const int* lens, float* hs, int T,
                         int B, int H, int ND, int reverse_mask, float cap, cudaStream_t st);
int rnn_relu_bwd_generic(const float* dout, const float* u, const int* lens, const float* hs, float* dz, int T, int B, int H,
                         int ND, int reverse_mask, float cap, cudaStream_t st);
int rnn_relu_fwd_persist(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs, float* hs32, int T, int B,
                         int H, int ND, int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st);
int rnn_relu_bwd_persist(const float* dout, const void* u, const void* ut, const int* lens, const void* hs, void* dz, float* dbias, int T, int B, int H, int ND,
                         int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st);
size_t rnn_persist_workspace_bytes(int T, int B, int H, int ND);
int rnn_persist_supported(int B, int H, int ND);

}  // namespace ds2

using namespace ds2;

extern "C" const char* ds2_last_error(void) { return g_err; }
extern "C" int ds2_version(void) { return 200; }
extern "C" void ds2_debug_reload_env(void) { (void)env(); reload_env(); }
extern "C" long long ds2_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

static int rnn_check(const void* a, const void* b, const void* c, const void* d, int T, int B, int H, int ND, int dtype) {
  DS2_REQUIRE(a && b && c && d, DS2_ERR_BAD_ARG, "ds2_rnn: null pointer");
  DS2_REQUIRE(T > 0 && B > 0 && H > 0 && (ND == 1 || ND == 2), DS2_ERR_BAD_SHAPE, "ds2_rnn: bad shape T=%d B=%d H=%d ND=%d", T, B, H, ND);
  DS2_REQUIRE(dtype == DS2_F32 || dtype == DS2_BF16, DS2_ERR_BAD_ARG, "ds2_rnn: bad dtype %d", dtype);
  return DS2_OK;
}

extern "C" int ds2_rnn_relu_fwd(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs,
                                int dtype, int T, int B, int H, int ND, int reverse_mask, float cap, void* workspace,
                                size_t workspace_bytes, void* stream) {
  int rc = rnn_check(wx, u, bias, hs, T, B, H, ND, dtype);
  if (rc) return rc;
  DS2_REQUIRE(lens, DS2_ERR_BAD_ARG, "ds2_rnn_relu_fwd: lens is null");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DS2_F32)
    return rnn_relu_fwd_generic(wx, scale, static_cast<const float*>(u), bias, lens, static_cast<float*>(hs), T, B, H, ND,
                                reverse_mask, cap, st);
  DS2_REQUIRE(rnn_persist_supported(B, H, ND), DS2_ERR_UNSUPPORTED,
              "ds2_rnn_relu_fwd: persistent bf16 kernel does not take B=%d H=%d ND=%d (use DS2_F32)", B, H, ND);
  return rnn_relu_fwd_persist(wx, scale, u, bias, lens, hs, nullptr, T, B, H, ND, reverse_mask, cap, workspace, workspace_bytes, st);
}

extern "C" int ds2_rnn_relu_fwd_ex(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs, float* hs_f32,
                                   int T, int B, int H, int ND, int reverse_mask, float cap, void* workspace, size_t workspace_bytes,
                                   void* stream) {
  int rc = rnn_check(wx, u, bias, hs, T, B, H, ND, DS2_BF16);
  if (rc) return rc;
  DS2_REQUIRE(lens, DS2_ERR_BAD_ARG, "ds2_rnn_relu_fwd_ex: lens is null");
  DS2_REQUIRE(rnn_persist_supported(B, H, ND), DS2_ERR_UNSUPPORTED,
              "ds2_rnn_relu_fwd_ex: persistent bf16 kernel does not take B=%d H=%d ND=%d", B, H, ND);
  return rnn_relu_fwd_persist(wx, scale, u, bias, lens, hs, hs_f32, T, B, H, ND, reverse_mask, cap, workspace, workspace_bytes,
                              static_cast<cudaStream_t>(stream));
}

extern "C" int ds2_rnn_relu_bwd(const float* dout, const void* u, const int* lens, const void* hs, void* dz, float* dbias,
                                int dtype, int T, int B, int H, int ND, int reverse_mask, float cap, void* workspace,
                                size_t workspace_bytes, void* stream) {
  int rc = rnn_check(dout, u, hs, dz, T, B, H, ND, dtype);
  if (rc) return rc;
  DS2_REQUIRE(lens, DS2_ERR_BAD_ARG, "ds2_rnn_relu_bwd: lens is null");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DS2_F32) {
    rc = rnn_relu_bwd_generic(dout, static_cast<const float*>(u), lens, static_cast<const float*>(hs), static_cast<float*>(dz),
                              T, B, H, ND, reverse_mask, cap, st);
    if (rc || !dbias) return rc;
    return ds2_colsum(dz, DS2_F32, (long)T * B, ND * H, (long)ND * H, dbias, 0, stream);
  }
  DS2_REQUIRE(rnn_persist_supported(B, H, ND), DS2_ERR_UNSUPPORTED,
              "ds2_rnn_relu_bwd: persistent bf16 kernel does not take B=%d H=%d ND=%d (use DS2_F32)", B, H, ND);
  return rnn_relu_bwd_persist(dout, u, nullptr, lens, hs, dz, dbias, T, B, H, ND, reverse_mask, cap, workspace, workspace_bytes, st);
}

extern "C" int ds2_rnn_relu_bwd_ex(const float* dout, const void* u, const void* ut, const int* lens, const void* hs, void* dz, float* dbias,
                                   int T, int B, int H, int ND, int reverse_mask, float cap, void* workspace, size_t workspace_bytes,
                                   void* stream) {
  int rc = rnn_check(dout, ut ? ut : u, hs, dz, T, B, H, ND, DS2_BF16);
  if (rc) return rc;
  DS2_REQUIRE(lens, DS2_ERR_BAD_ARG, "ds2_rnn_relu_bwd_ex: lens is null");
  DS2_REQUIRE(rnn_persist_supported(B, H, ND), DS2_ERR_UNSUPPORTED,
              "ds2_rnn_relu_bwd_ex: persistent bf16 kernel does not take B=%d H=%d ND=%d", B, H, ND);
  return rnn_relu_bwd_persist(dout, u, ut, lens, hs, dz, dbias, T, B, H, ND, reverse_mask, cap, workspace, workspace_bytes,
                              static_cast<cudaStream_t>(stream));
}

extern "C" size_t ds2_rnn_workspace_bytes(int T, int B, int H, int ND, int dtype) {
  if (dtype != DS2_BF16) return 0;
  return rnn_persist_workspace_bytes(T, B, H, ND);
}
extern "C" int ds2_set_sm_limit(int n) { return set_sm_limit(n); }

extern "C" int ds2_rnn_relu_persistent_supported(int B, int H, int ND) { return rnn_persist_supported(B, H, ND); }
