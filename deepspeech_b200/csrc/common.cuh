// Shared device/host helpers for the ds2_b200 kernels (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ds2_b200.h"

namespace ds2 {

// ------------------------------------------------------------------ error plumbing (host)
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
#define DS2_CUDA(call)                                             \
  do {                                                             \
    cudaError_t _e = (call);                                       \
    if (_e != cudaSuccess) return ::ds2::cuda_fail(_e, #call);     \
  } while (0)
#define DS2_LAUNCH_CHECK(name)                                     \
  do {                                                             \
    cudaError_t _e = cudaGetLastError();                           \
    if (_e != cudaSuccess) return ::ds2::cuda_fail(_e, name);      \
  } while (0)
#define DS2_REQUIRE(cond, code, ...)                               \
  do {                                                             \
    if (!(cond)) {                                                 \
      ::ds2::set_error(__VA_ARGS__);                               \
      return (code);                                               \
    }                                                              \
  } while (0)

// Environment switches (debugging / A-B runs), read ONCE per process (ds2_debug_reload_env re-reads them).
struct EnvSwitches {
  int rnn_no_coop;      // DS2_RNN_NO_COOP: never use the cooperative launch attribute
  int rnn_exp;          // DS2_RNN_EXP: timing experiments (results invalid for some bits)
  int rnn_fenced;       // DS2_RNN_FENCED: release/acquire step counters instead of self-validating rows
  int rnn_slab_f32;     // DS2_RNN_SLAB_F32: fp32 partial sums across the cluster
  int rnn_prof;         // DS2_RNN_PROF: in-kernel stage stamps
  int rnn_v1;           // DS2_RNN_V1: round-1 exchange protocol of the ReLU recurrence (A/B runs)
  int rnn_poll_margin;  // DS2_RNN_POLL_MARGIN: clk before a step's predicted completion at which the counter polling starts (default: per batch size, 725-900; 0: always poll)
  int gru_no_bulk;      // DS2_GRU_NO_BULK
  int gemm_no_pair;     // DS2_GEMM_NO_PAIR
  int ctc_ns;           // DS2_CTC_NS
  int ctc_v1;           // DS2_CTC_V1: round-1 CTC kernels (A/B runs)
  int profiler;         // a profiler / injection library is attached to this process (ncu, nsys, compute-sanitizer)
};
const EnvSwitches& env();
void reload_env();
// Cooperative launches guarantee co-residency of the persistent grids but cannot be replayed by profilers:
// used unless DS2_RNN_NO_COOP is set or a profiler is attached.
bool use_cooperative_launch();

int num_sms();              // SM count of the current device, capped by ds2_set_sm_limit
int set_sm_limit(int n);
// true the first time it is called with `flags` on the current device (kernel attributes such as the dynamic
// shared-memory limit are per device: a process that drives several GPUs must set them on each)
bool first_use_on_device(unsigned long long* flags);
void count_launch(int n = 1);

// ------------------------------------------------------------------ small device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// One lane of a fully converged warp (the lowest one).  ptxas knows an elect.sync-guarded region has a single
// active thread and feeds the uniform-datapath instructions (UTCHMMA / UTMALDG / UTCBAR) directly; a plain
// `lane == 0` branch makes it emit a vote + R2UR.BROADCAST waterfall loop around every such instruction.
__device__ __forceinline__ bool elect_one_sync() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
// warp index as a value the compiler can prove warp-uniform
__device__ __forceinline__ int uniform_warp_idx() { return __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0); }
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// Device-side watchdog: every spin loop in the library is bounded so that a protocol bug
// traps (cudaErrorLaunchFailure) instead of hanging the GPU.
#ifndef DS2_SPIN_LIMIT_CYCLES
#define DS2_SPIN_LIMIT_CYCLES (4000000000ll)   // ~2 s at 1.9 GHz
#endif
__device__ __forceinline__ long long dev_clock() {
#ifdef __CUDA_ARCH__
  return clock64();
#else
  return 0;
#endif
}
// Only the first few waiters report (one line each, with the tag of the wait they were stuck in); the others trap
// silently - a timed-out grid has thousands of stuck threads and their lines would bury the first, informative one.
static __device__ unsigned int g_spin_reports = 0;
struct SpinGuard {
  long long t0;
  __device__ __forceinline__ SpinGuard() : t0(dev_clock()) {}
  __device__ __forceinline__ void check(int tag) {
    if (dev_clock() - t0 > DS2_SPIN_LIMIT_CYCLES) {
      if (atomicAdd(&g_spin_reports, 1u) < 4u)
        printf("ds2: device spin timeout tag=%d block=(%d,%d,%d) thread=%d\n", tag, blockIdx.x, blockIdx.y, blockIdx.z, threadIdx.x);
      __trap();
    }
  }
};

// Explicit shared-space accesses by 32-bit shared address.  Pointers derived from an integer-aligned dynamic
// shared base lose their address space, and the compiler then emits GENERIC ld/st (a dozen address instructions
// each and a slower path through the LSU); these keep hot-loop staging on STS/LDS.
__device__ __forceinline__ void sts_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

// ------------------------------------------------------------------ fence-free publication of bf16 rows
// The persistent recurrence kernels hand each step's rows to the other CTAs through global memory.  A release
// fence before the step-counter increment costs ~1000 cycles per step (MEMBAR.GPU waits for every outstanding
// store of the SM), so the increment is RELAXED and the data validates itself instead: the host fills the exchanged
// buffer with 0xFF bytes before the launch, every location is written exactly once per launch, writers never emit
// the bf16 pattern 0xFFFF, and a reader that still sees the fill after the counter said "ready" simply reloads.
// Each writer stores 8 aligned bytes (4 bf16) at once, so one bf16 per 8-byte half tells whether the half landed.
static constexpr int DS2_FILL_BYTE = 0xFF;
__device__ __forceinline__ bool unit_stale(const uint4& v) {
  return ((v.x & 0xFFFFu) == 0xFFFFu) || ((v.z & 0xFFFFu) == 0xFFFFu);
}
// a negative NaN with all payload bits set is the only bf16 value equal to the fill: make it the positive one
__device__ __forceinline__ uint32_t scrub_fill(uint32_t w) {
  if ((w & 0xFFFFu) == 0xFFFFu) w &= 0xFFFF7FFFu;
  if ((w >> 16) == 0xFFFFu) w &= 0x7FFFFFFFu;
  return w;
}
__device__ __forceinline__ unsigned int ld_relaxed_gpu_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_relaxed_gpu_add(unsigned int* p, unsigned int v) {
  asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ------------------------------------------------------------------ mbarrier / TMA / tcgen05 PTX
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// The uint32_t overloads take a shared-space address computed once (see keep_reg): hot loops use them so that no
// generic->shared conversion (two special-register reads) sits in front of every barrier operation.
__device__ __forceinline__ uint32_t keep_reg(uint32_t x) {
  asm volatile("" : "+r"(x));      // opaque to the optimiser: the value stays in a register, it is not recomputed
  return x;
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) { mbar_arrive_expect_tx(smem_u32(bar), bytes); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) { mbar_arrive(smem_u32(bar)); }
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) { return mbar_try_wait(smem_u32(bar), parity); }
// non-blocking form (try_wait may suspend the thread for a hardware-defined time): for loops that also poll other things
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int tag) {
  if (mbar_try_wait(bar, parity)) return;
  SpinGuard g;
  while (!mbar_try_wait(bar, parity)) g.check(tag);
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int tag) { mbar_wait(smem_u32(bar), parity, tag); }

__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// TMEM allocation: called by ONE full warp.  ncols: power of two in [32,512].
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 x bf16 -> f32, issued by one thread.
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed.
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) { umma_commit(smem_u32(bar)); }
// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns (thread i <- lane base+i).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor for tcgen05.mma, SWIZZLE_128B canonical layouts
// (bit layout: cute::UMMA::SmemDescriptor; start/LBO/SBO in 16-byte units, version=1, layout=2).
__device__ __forceinline__ uint64_t umma_smem_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;   // descriptor version (Blackwell)
  d |= static_cast<uint64_t>(2) << 61;   // SWIZZLE_128B
  return d;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): bf16 x bf16 -> f32.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(a_mn_major) << 15) |
         (static_cast<uint32_t>(b_mn_major) << 16) | (static_cast<uint32_t>(N >> 3) << 17) |
         (static_cast<uint32_t>(M >> 4) << 24);
}

// ------------------------------------------------------------------ host-side TMA descriptor factory
// 2-D bf16 tensor map, 128B swizzle.  dims = {inner, outer} elements, stride_outer in elements.
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t stride_outer_elems,
                      uint32_t box_inner, uint32_t box_outer);

}  // namespace ds2
