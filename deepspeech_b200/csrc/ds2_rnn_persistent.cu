// Persistent, weights-resident recurrent kernel (clipped-ReLU RNN), forward and backward.
//
// Replaces the tf.while_loop of tf.nn.bidirectional_dynamic_rnn around CustomRNNCell2
// (src/custom_ops.py:48-72, :91-93): per timestep  h_t = clip(relu(s*Wx_t + U h_{t-1} + b)).
// The per-step product U.h is a [H x H] x [H x B] contraction on the tensor cores:
//   * the H output units are tiled by 128 (tcgen05 M=128), the batch is the MMA N dimension;
//   * the reduction dimension is split over a thread-block cluster of 4 CTAs; each CTA loads its
//     [128 x H/4] bf16 slice of U ONCE (TMA -> shared memory -> tcgen05.cp -> TMEM) and keeps it resident in
//     tensor memory as the MMA's A operand for all T steps;
//   * the batch is cut into independent 16-utterance chains ("instances") that share the resident U: each
//     instance has its own control warp, 4 loader/epilogue warps, barriers and accumulator columns, so one
//     chain's memory latency is covered by the other chains' work;
//   * timesteps are ordered per (direction, instance) by a release/acquire step counter in global memory:
//     one thread per instance acquires it, then the instance's 128 loader threads fetch their K-slice of
//     h_{t-1} with ordinary 16-byte loads and write it into shared memory in the 128B-swizzled K-major layout
//     the MMA descriptor expects (a TMA of the 7 thin [16 x 64] boxes measured ~2500 clk + a 700 clk
//     cross-proxy fence; the loads + a shared-memory-only proxy fence take a fraction of that);
//   * each step of an instance: acquire -> load h_{t-1} slice -> tcgen05.mma into TMEM -> tcgen05.ld ->
//     partial sums pushed with st.async (completes a transaction barrier in the receiving CTA, no fence) to
//     the cluster CTA that finishes those 32 rows -> epilogue (add s*Wx_t + b, clip, length mask) -> bf16
//     h_t to global -> release-increment of the counter;
//   * both directions run concurrently as independent groups of CTAs of the same launch.
// The backward kernel is the same machine with A = U^T, moving operand dz_{t+1}, and the epilogue
// dz_t = [0 < h_t < cap] * (dout_t + U^T dz_{t+1}).
// Every spin loop is bounded by the device watchdog (common.cuh), so a protocol bug traps instead of hanging.
#include <stdlib.h>

#include "common.cuh"

namespace ds2 {

static constexpr int KS = 4;              // cluster size = K split
static constexpr int BP = 16;             // utterances per instance (MMA N)
static constexpr int INST_THREADS = 160;  // per instance: 1 control warp + 4 loader/epilogue warps
static constexpr int MAX_NI = 4;
static constexpr int MAX_NBOX = 9;        // K-slice <= 576 columns
static constexpr int MAX_UNITS = (BP * MAX_NBOX * 8 + 127) / 128;   // 16-byte units per loader thread

struct RnnPersistParams {
  const float* add_in;     // fwd: wx [T,B,ND*H] ; bwd: dout [T,B,H]
  const float* bias;       // fwd: [ND,H]
  const int* lens;         // [B]
  const __nv_bfloat16* hs_in;   // bwd: hs [ND,T,B,H] (activation mask)
  float* dbias;            // bwd, optional: [ND,H] receives sum_t,b dz (zero-filled by the host side)
  __nv_bfloat16* out;      // fwd: hs [ND,T,B,H] ; bwd: dz [T,B,ND*H]   (also the moving operand of the next step)
  unsigned int* ctr;       // [ND][NI] step counters, 32 uints apart
  long long* prof;         // optional per-step clock64 stamps of CTA (0,0,0) instance 0 (DS2_RNN_PROF=1), 16 per step
  long out_row_pitch;      // elements between consecutive (t,b) rows of `out`
  long out_dir_stride;     // element offset of direction d inside `out`
  int T, B, H, ND, reverse_mask;
  int NI;                  // instances = B / BP
  int tiles_m, ksl, nbox;  // M tiles per direction, K-slice length (multiple of 64), boxes per slice
  float scale, cap;
  int exp;                 // timing experiments (DS2_RNN_EXP, results are NOT valid when set)
  int slab16;              // 1: partial sums cross the cluster as bf16 (default), 0: fp32 (DS2_RNN_SLAB_F32=1)
  int fenced;              // 1: release/acquire step counters (DS2_RNN_FENCED=1); 0: relaxed counters + self-validating
                           // rows (see common.cuh, `out` must have been filled with DS2_FILL_BYTE)
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t map_to_rank(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
// asynchronous 16-byte store into a cluster CTA's shared memory; completes 16 bytes on that CTA's mbarrier
__device__ __forceinline__ void st_async_v4(uint32_t remote_addr, uint32_t remote_bar, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%2, %3, %4, %5}, [%1];"
               ::"r"(remote_addr), "r"(remote_bar), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 ld_relaxed_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_cg_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// smem (matrix descriptor, same format as the MMA operand descriptors) -> TMEM, 128 lanes x 256 bits
__device__ __forceinline__ void tmem_cp_128x256b(uint32_t taddr, uint64_t sdesc) {
  asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// shared-memory carve-up (all offsets from a 1024-aligned base)
struct RnnSmem {
  uint32_t a_off, b_off, slab_off, bar_off, total;
};
static constexpr int BARS_PER_INST = 4;   // b_ready, mma_done, slab_full[2]
__host__ __device__ inline RnnSmem rnn_smem_layout(int nbox, int NI) {
  RnnSmem s;
  s.a_off = 0;                                            // nbox x [128 x 64] bf16 (SW128 K-major), staging for TMEM
  s.b_off = s.a_off + nbox * 16384;                       // NI x nbox x [16 x 64] bf16
  s.slab_off = s.b_off + NI * nbox * (BP * 128);          // NI x KS x [32 x 16] fp32
  s.bar_off = s.slab_off + NI * KS * 32 * BP * 4;         // mbarriers
  s.total = s.bar_off + 8 * (2 + NI * BARS_PER_INST) + 16;
  return s;
}

// NIMAX bounds the block size (INST_THREADS * NI threads) and with it the registers a thread may use: the common
// two-instance launch (B = 32) gets 168 registers, the four-instance one (B = 64) has to live with 96.
template <bool BWD, int NIMAX>
__global__ void __launch_bounds__(INST_THREADS * NIMAX, 1)
rnn_persist_kernel(const __grid_constant__ CUtensorMap tm_a, RnnPersistParams p) {
  extern __shared__ uint8_t rnn_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(rnn_smem_raw) + 1023) & ~uintptr_t(1023));
  const int NI = p.NI, nbox = p.nbox;
  const RnnSmem L = rnn_smem_layout(nbox, NI);
  uint8_t* sA = smem + L.a_off;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
  uint64_t* a_bar = bars;            // U slice landed in smem
  uint64_t* acp_bar = bars + 1;      // U slice copied into TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 + NI * BARS_PER_INST);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int inst = warp / 5, w_in = warp % 5;
  uint64_t* b_ready = bars + 2 + inst * BARS_PER_INST;      // moving operand of this step is in shared memory
  uint64_t* mma_bar = b_ready + 1;                          // accumulators ready
  uint64_t* slab_bar = b_ready + 2;                         // partial slabs of this step arrived (tx bytes); a sender
                                                            // cannot be a step ahead: it first acquires the step counter
  uint8_t* sB = smem + L.b_off + inst * nbox * (BP * 128);
  float* slab = reinterpret_cast<float*>(smem + L.slab_off) + inst * KS * 32 * BP;
  // shared-space addresses used inside the step loop, computed once
  const uint32_t b_ready_a = keep_reg(smem_u32(b_ready)), mma_bar_a = b_ready_a + 8, slab_bar_a = b_ready_a + 16;
  const uint32_t sB_addr = keep_reg(smem_u32(sB)), slab_local = keep_reg(smem_u32(slab));

  const uint32_t rank = cluster_ctarank();                  // K slice owned by this CTA, rows it finishes
  const int mt = blockIdx.y, d = blockIdx.z;
  const int m0 = mt * 128;
  const int rev = (p.reverse_mask >> d) & 1;
  const int T = p.T, H = p.H, ND = p.ND, Bn = p.B;
  unsigned int* ctr = p.ctr + (d * NI + inst) * 32;
  const unsigned int ctas_per_dir = KS * p.tiles_m;
  const bool prof_cta = p.prof && (blockIdx.x + blockIdx.y + blockIdx.z == 0) && inst == 0;
  const int acols = nbox * 32;                              // TMEM columns holding U (8 per K=16 chunk)

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm_a);
    mbar_init(a_bar, 1);
    mbar_init(acp_bar, 1);
    for (int i = 0; i < NI; ++i) {
      uint64_t* bb = bars + 2 + i * BARS_PER_INST;
      mbar_init(bb, 128);          // every loader thread arrives once per step
      mbar_init(bb + 1, 1);        // tcgen05.commit
      mbar_init(bb + 2, 1);        // one arrive.expect_tx per use + the st.async bytes
      mbar_init(bb + 3, 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t d_tmem = tmem_base + acols + inst * BP;
  // every CTA of the cluster must have initialised its barriers before any remote st.async
  cluster_arrive();
  cluster_wait();

  if (w_in == 0) {
    // =========================== control warp of this instance: MMA issue ===========================
    if (inst == 0 && elect_one_sync()) {
      mbar_arrive_expect_tx(a_bar, nbox * 16384);
      for (int i = 0; i < nbox; ++i) tma_load_2d(sA + i * 16384, &tm_a, a_bar, rank * p.ksl + i * 64, d * H + m0);
      mbar_wait(a_bar, 0, 10);
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA);
      for (int c = 0; c < nbox * 4; ++c)
        tmem_cp_128x256b(tmem_base + c * 8, umma_smem_desc_sw128(a0 + (c >> 2) * 16384 + (c & 3) * 32, 16, 1024));
      umma_commit(acp_bar);          // arrives when the copies have completed
    }
    __syncwarp();
    mbar_wait(acp_bar, 0, 11);       // every instance's control warp: U is resident in TMEM
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16(128, BP, 0, 0);
    const uint32_t b0 = sB_addr;
    uint32_t ph = 0;
    for (int step = 1; step < T; ++step) {
      mbar_wait(b_ready_a, ph, 30);    // loaders have written h_{t-1} (generic proxy) and fenced it to the async proxy
      tc_fence_after();
      if (elect_one_sync()) {
        if (prof_cta && step < 1024) p.prof[step * 16 + 0] = dev_clock();
        for (int i = 0; i < nbox; ++i) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16_ts(d_tmem, tmem_base + (i * 4 + k) * 8, umma_smem_desc_sw128(b0 + i * (BP * 128) + k * 32, 16, 1024), idesc,
                         (i > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(mma_bar_a);
        if (prof_cta && step < 1024) p.prof[step * 16 + 1] = dev_clock();
      }
      __syncwarp();
      ph ^= 1;
    }
  } else {
    // =========================== loader + epilogue warps of this instance ===========================
    const int q = warp & 3;                      // TMEM lane quarter = rows [32q, 32q+32) of the tile
    const int et = (w_in - 1) * 32 + lane;       // 0..127
    const int bl = et % BP;                      // utterance within the instance
    const int jg = et / BP;                      // 0..7: 4-row group of this CTA's 32 rows
    const int b = inst * BP + bl;
    const int j = m0 + 32 * (int)rank + 4 * jg;  // first of 4 consecutive output units
    const bool jvalid = j < H;                   // H % 8 == 0 -> all-or-nothing
    float bias4[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) bias4[i] = (!BWD && jvalid) ? p.bias[d * H + j + i] : 0.f;
    const int len_b = p.lens[b];
    const uint32_t slab_bar_local = slab_bar_a;
    // loader role: 16-byte units (8 consecutive k of one utterance) of this CTA's K slice
    const int cpr = p.ksl / 8;                   // units per utterance row
    const int nunits = BP * cpr;
    const __nv_bfloat16* src_dir = p.out + (size_t)d * p.out_dir_stride;
    // per-thread unit table, computed once: global element offset inside the step's row block (-1: no load,
    // the unit lies beyond H and is zero-filled) and swizzled byte offset inside this instance's B tile
    int goff[MAX_UNITS];
    int soff[MAX_UNITS];
#pragma unroll
    for (int m = 0; m < MAX_UNITS; ++m) {
      const int u = et + 128 * m;
      goff[m] = -1;
      soff[m] = -1;
      if (u < nunits) {
        const int ub = u / cpr, uc = u - ub * cpr;
        // box (uc/8) of [16 rows x 128 B], 16-byte chunk (uc%8) XOR-swizzled with the row (SWIZZLE_128B)
        soff[m] = (uc >> 3) * (BP * 128) + ub * 128 + (((uc & 7) ^ (ub & 7)) << 4);
        if ((int)rank * p.ksl + uc * 8 < H) goff[m] = ub * (int)p.out_row_pitch + uc * 8;
      }
    }
    if ((p.exp & 16) && (inst & 1)) { __nanosleep(900); __nanosleep(900); }
    if ((p.exp & 32) && d == 1) __nanosleep(900);
    float dbs[4] = {0.f, 0.f, 0.f, 0.f};         // backward: running sum over time of this thread's dz (bias gradient)
    uint32_t ph = 0;
    for (int step = 0; step < T; ++step) {
      const int t = BWD ? (rev ? step : T - 1 - step) : (rev ? T - 1 - step : step);
      const int tsrc = BWD ? (rev ? t - 1 : t + 1) : (rev ? t + 1 : t - 1);   // step whose result feeds this one
      const bool pr = prof_cta && et == 0 && step < 1024;
      // prefetch the additive term (s*Wx_t or dout_t) and, for the backward, h_t for the mask
      float4 add = make_float4(0.f, 0.f, 0.f, 0.f);
      uint2 hm = make_uint2(0u, 0u);
      if (jvalid) {
        const float* src = BWD ? p.add_in + ((size_t)t * Bn + b) * H + j
                               : p.add_in + ((size_t)t * Bn + b) * ND * H + (size_t)d * H + j;
        add = __ldg(reinterpret_cast<const float4*>(src));
        if (BWD) hm = *reinterpret_cast<const uint2*>(p.hs_in + (((size_t)d * T + t) * Bn + b) * H + j);
      }
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      if (step > 0) {
        if (et == 0) mbar_arrive_expect_tx(slab_bar_a, p.slab16 ? KS * 32 * BP * 2 : KS * 32 * BP * 4);   // arm this step's slab barrier
        // ---- loader: poll h_{t-1}[inst's utterances][this CTA's K slice] until written, stage into smem
        if (pr) p.prof[step * 16 + 8] = dev_clock();
        {
          const __nv_bfloat16* rows = src_dir + ((size_t)tsrc * Bn + inst * BP) * p.out_row_pitch + (size_t)rank * p.ksl;
          // one thread acquires "every CTA of this direction has published step-1 of this instance";
          // the named barrier hands that visibility to the other 127 loader threads
          if (et == 0) {
            const unsigned int need = ctas_per_dir * (unsigned int)step;
            if (p.fenced) {
              if (!(p.exp & 2) && ld_acquire_u32(ctr) < need) {
                SpinGuard g;
                while (ld_acquire_u32(ctr) < need) g.check(20);
              }
            } else if (!(p.exp & 2) && ld_relaxed_gpu_u32(ctr) < need) {
              SpinGuard g;
              while (ld_relaxed_gpu_u32(ctr) < need) g.check(21);
            }
          }
          asm volatile("bar.sync %0, 128;" ::"r"(1 + inst) : "memory");
          if (pr) p.prof[step * 16 + 15] = dev_clock();
          uint4 v[MAX_UNITS];
#pragma unroll
          for (int m = 0; m < MAX_UNITS; ++m) {
            v[m] = make_uint4(0u, 0u, 0u, 0u);
            if (goff[m] >= 0 && !(p.exp & 4)) v[m] = (p.exp & 256) ? ld_cg_v4(rows + goff[m]) : ld_relaxed_v4(rows + goff[m]);
          }
          if (pr) p.prof[step * 16 + 9] = dev_clock();
          int n_retry = 0;
          if (!p.fenced && !(p.exp & 6)) {
            // the counter was published without a fence: a row may still show the host's fill; reload until it landed
#pragma unroll
            for (int m = 0; m < MAX_UNITS; ++m) {
              if (goff[m] >= 0 && unit_stale(v[m])) {
                SpinGuard g;
                do {
                  v[m] = ld_relaxed_v4(rows + goff[m]);
                  g.check(22);
                  ++n_retry;
                } while (unit_stale(v[m]));
              }
            }
          }
#pragma unroll
          for (int m = 0; m < MAX_UNITS; ++m)
            if (soff[m] >= 0) sts_v4(sB_addr + soff[m], v[m]);
          if (pr) { p.prof[step * 16 + 13] = dev_clock(); p.prof[step * 16 + 2] = n_retry; }
          fence_proxy_async();       // generic-proxy smem writes -> visible to the tensor core (async proxy)
          if (pr) p.prof[step * 16 + 14] = dev_clock();
          mbar_arrive(b_ready_a);
        }
        if (pr) p.prof[step * 16 + 10] = dev_clock();
        // ---- epilogue part 1: partial sums of this CTA's K slice -> the CTA that finishes those rows
        mbar_wait(mma_bar_a, ph, 40);
        tc_fence_after();
        if (pr) p.prof[step * 16 + 3] = dev_clock();
        uint32_t r[16];
        tmem_ld_32x16(d_tmem + (static_cast<uint32_t>(q * 32) << 16), r);
        tmem_ld_wait();
        tc_fence_before();
        if (pr) p.prof[step * 16 + 11] = dev_clock();
        {
          const uint32_t off = ((rank * 32 + lane) * BP) * 4;
          const uint32_t dst = map_to_rank(slab_local, (uint32_t)q) + off;
          const uint32_t rbar = map_to_rank(slab_bar_local, (uint32_t)q);
          if (p.exp & 128) {
            // experiment: the warp parks its 32 rows (2 KB) in local smem and ONE bulk copy carries them to the finisher
            const uint32_t stg = smem_u32(sA) + (inst * 4 + (w_in - 1)) * 2048;
#pragma unroll
            for (int c = 0; c < 4; ++c) sts_v4(stg + lane * 64 + c * 16, make_uint4(r[4 * c], r[4 * c + 1], r[4 * c + 2], r[4 * c + 3]));
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              const uint32_t dblk = map_to_rank(slab_local, (uint32_t)q) + rank * 32 * BP * 4;
              asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                           ::"r"(dblk), "r"(stg), "r"(2048), "r"(rbar) : "memory");
            }
          } else if (p.slab16) {
            // The exchange is bound by the DSMEM store queue (4 st.async per thread: 690 clk to issue, 3.76 us/step);
            // as bf16 it is 2 per thread (110 clk, 3.26 us/step).  The rounding adds ~10 % to the error the bf16
            // storage of h makes anyway (tools/rnn_err.py: rms 1.65e-3 -> 1.83e-3 of the output scale).
            uint32_t h[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(r[2 * c]), __uint_as_float(r[2 * c + 1]));
              h[c] = *reinterpret_cast<uint32_t*>(&t2);
            }
            const uint32_t dst2 = map_to_rank(slab_local, (uint32_t)q) + off / 2;
            st_async_v4(dst2, rbar, h[0], h[1], h[2], h[3]);
            st_async_v4(dst2 + 16, rbar, h[4], h[5], h[6], h[7]);
          } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) st_async_v4(dst + c * 16, rbar, r[4 * c], r[4 * c + 1], r[4 * c + 2], r[4 * c + 3]);
          }
        }
        if (pr) p.prof[step * 16 + 12] = dev_clock();
        // ---- epilogue part 2: wait for the KS slabs of this step, reduce
        mbar_wait(slab_bar_a, ph, 50);
        if (pr) p.prof[step * 16 + 5] = dev_clock();
#pragma unroll
        for (int s = 0; s < KS; ++s) {
          const uint32_t sp = slab_local + ((s * 32 + 4 * jg) * BP + bl) * 4;
          if (p.slab16) {
            const uint32_t sp2 = slab_local + ((s * 32 + 4 * jg) * BP + bl) * 2;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              unsigned short hv;
              asm volatile("ld.shared.u16 %0, [%1];" : "=h"(hv) : "r"(sp2 + i * BP * 2) : "memory");
              acc[i] += __uint_as_float((uint32_t)hv << 16);
            }
          } else {
            acc[0] += lds_f32(sp); acc[1] += lds_f32(sp + BP * 4); acc[2] += lds_f32(sp + 2 * BP * 4); acc[3] += lds_f32(sp + 3 * BP * 4);
          }
        }
        ph ^= 1;
      }
      if (jvalid) {
        const bool live = t < len_b;
        const float av[4] = {add.x, add.y, add.z, add.w};
        float o[4];
        if (!BWD) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float z = fmaf(p.scale, av[i], acc[i] + bias4[i]);
            o[i] = live ? fminf(fmaxf(z, 0.f), p.cap) : 0.f;
          }
        } else {
          const __nv_bfloat16* hb = reinterpret_cast<const __nv_bfloat16*>(&hm);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float h = __bfloat162float(hb[i]);
            o[i] = (live && h > 0.f && h < p.cap) ? (av[i] + acc[i]) : 0.f;
            dbs[i] += o[i];
          }
        }
        uint2 pk;
        __nv_bfloat162 b0v = __floats2bfloat162_rn(o[0], o[1]), b1v = __floats2bfloat162_rn(o[2], o[3]);
        pk.x = scrub_fill(*reinterpret_cast<uint32_t*>(&b0v));
        pk.y = scrub_fill(*reinterpret_cast<uint32_t*>(&b1v));
        __nv_bfloat16* dstp = p.out + (size_t)d * p.out_dir_stride + ((size_t)t * Bn + b) * p.out_row_pitch + j;
        *reinterpret_cast<uint2*>(dstp) = pk;
      }
      if (pr) p.prof[step * 16 + 6] = dev_clock();
      // publish: the instance's 128 threads' stores -> one release increment of its step counter
      // (the CTA-level barrier orders their stores before thread 0's gpu-scope release: cumulativity)
      if (!(p.exp & 8)) asm volatile("bar.sync %0, 128;" ::"r"(1 + inst) : "memory");
      if (et == 0) {
        if (p.fenced) red_release_add(ctr, 1u);
        else red_relaxed_gpu_add(ctr, 1u);
      }
      if (pr) p.prof[step * 16 + 7] = dev_clock();
    }
    if (BWD && p.dbias != nullptr) {
      // bias gradient: sum over the instance's 16 utterances (16 consecutive lanes share a unit quad), one red per unit
#pragma unroll
      for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) dbs[i] += __shfl_xor_sync(0xffffffffu, dbs[i], o);
      }
      if (bl == 0 && jvalid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) atomicAdd(p.dbias + (size_t)d * H + j + i, dbs[i]);
      }
    }
  }
  // no CTA may exit while a peer can still write into its shared memory
  tc_fence_before();
  cluster_arrive();
  cluster_wait();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------ host side
static int persist_geometry(int B, int H, int ND, int* tiles_m, int* ksl, int* nbox, int* ni, size_t* smem) {
  if (B % BP != 0 || B < BP || B / BP > MAX_NI || H % 8 != 0 || H < 8 || (ND != 1 && ND != 2)) return 0;
  const int tm = (H + 127) / 128;
  const int slice = (H + KS - 1) / KS;
  const int l = ((slice + 63) / 64) * 64;
  const int nb = l / 64;
  if (nb > MAX_NBOX) return 0;
  const int n_inst = B / BP;
  if (nb * 32 + n_inst * BP > 512) return 0;   // TMEM columns: resident U + accumulators
  const size_t bytes = rnn_smem_layout(nb, n_inst).total + 1024;
  if (bytes > 225 * 1024) return 0;
  if (KS * tm * ND > 128) return 0;           // must be co-resident in clusters of 4 (148 SMs, 1 CTA / SM)
  if (tiles_m) *tiles_m = tm;
  if (ksl) *ksl = l;
  if (nbox) *nbox = nb;
  if (ni) *ni = n_inst;
  if (smem) *smem = bytes;
  return 1;
}


static constexpr size_t WS_CTR_BYTES = 2048;     // ND x MAX_NI counters, one 128-byte line each
static constexpr size_t WS_PROF_BYTES = 131072;

size_t rnn_persist_workspace_bytes_v1(int T, int B, int H, int ND) {
  (void)T; (void)B;
  // optional profile stamps + room for U^T (bf16) used by the backward
  return WS_CTR_BYTES + WS_PROF_BYTES + (size_t)ND * H * H * 2;
}

template <bool BWD, int NIMAX>
static int persist_launch_n(const CUtensorMap& ta, RnnPersistParams& p, size_t smem, cudaStream_t st) {
  auto kern = rnn_persist_kernel<BWD, NIMAX>;
  static unsigned long long attr_done_devs = 0;
  if (first_use_on_device(&attr_done_devs)) {
    DS2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(KS, p.tiles_m, p.ND);
  cfg.blockDim = dim3(INST_THREADS * p.NI);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = KS;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;       // all CTAs co-resident (they wait on each other's data)
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = use_cooperative_launch() ? 2 : 1;   // profilers cannot replay cooperative launches
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, p);
  if (e != cudaSuccess && cfg.numAttrs == 2) {
    // some driver versions reject cooperative + cluster launches: the grid (<= 128 CTAs, 1 per SM)
    // fits the device anyway, so fall back to a plain cluster launch
    (void)cudaGetLastError();
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, kern, ta, p);
  }
  if (e != cudaSuccess) return cuda_fail(e, BWD ? "rnn_persist_kernel<bwd>" : "rnn_persist_kernel<fwd>");
  count_launch();
  return DS2_OK;
}

template <bool BWD>
static int persist_launch(const CUtensorMap& ta, RnnPersistParams& p, size_t smem, cudaStream_t st) {
  return p.NI <= 2 ? persist_launch_n<BWD, 2>(ta, p, smem, st) : persist_launch_n<BWD, MAX_NI>(ta, p, smem, st);
}

static int persist_common(RnnPersistParams& p, size_t* smem, int T, int B, int H, int ND, void* ws, size_t ws_bytes, const char* who,
                          cudaStream_t st) {
  DS2_REQUIRE(persist_geometry(B, H, ND, &p.tiles_m, &p.ksl, &p.nbox, &p.NI, smem), DS2_ERR_UNSUPPORTED,
              "%s: persistent kernel does not take B=%d H=%d ND=%d", who, B, H, ND);
  const size_t need = rnn_persist_workspace_bytes_v1(T, B, H, ND);
  DS2_REQUIRE(ws && ws_bytes >= need, DS2_ERR_WORKSPACE, "%s: workspace of %zu bytes required", who, need);
  DS2_CUDA(cudaMemsetAsync(ws, 0, WS_CTR_BYTES, st));
  p.ctr = static_cast<unsigned int*>(ws);
  p.exp = env().rnn_exp;
  p.fenced = env().rnn_fenced;
  p.slab16 = (env().rnn_slab_f32 || (p.exp & 128)) ? 0 : 1;
  p.prof = env().rnn_prof ? reinterpret_cast<long long*>(static_cast<char*>(ws) + WS_CTR_BYTES) : nullptr;
  return DS2_OK;
}

int rnn_relu_fwd_persist_v1(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs, int T, int B,
                         int H, int ND, int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st) {
  RnnPersistParams p = {};
  size_t smem;
  int rc = persist_common(p, &smem, T, B, H, ND, ws, ws_bytes, "ds2_rnn_relu_fwd", st);
  if (rc) return rc;
  CUtensorMap ta;
  rc = make_tmap_2d_bf16(&ta, u, H, (uint64_t)ND * H, H, 64, 128);
  if (rc) return rc;
  p.add_in = wx; p.bias = bias; p.lens = lens; p.hs_in = nullptr; p.out = static_cast<__nv_bfloat16*>(hs);
  p.out_row_pitch = H; p.out_dir_stride = (long)T * B * H;
  p.T = T; p.B = B; p.H = H; p.ND = ND; p.reverse_mask = reverse_mask; p.scale = scale; p.cap = cap;
  if (!p.fenced) DS2_CUDA(cudaMemsetAsync(hs, DS2_FILL_BYTE, (size_t)ND * T * B * H * 2, st));
  return persist_launch<false>(ta, p, smem, st);
}

// The backward contraction is U^T dz: U is transposed per direction into the workspace first
// (ut[d][k][j] = U[d][j][k], 2*ND*H*H bytes, one small kernel per direction), then used as the resident A operand.
int rnn_relu_bwd_persist_v1(const float* dout, const void* u, const int* lens, const void* hs, void* dz, float* dbias, int T, int B, int H,
                         int ND, int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st) {
  RnnPersistParams p = {};
  size_t smem;
  int rc = persist_common(p, &smem, T, B, H, ND, ws, ws_bytes, "ds2_rnn_relu_bwd", st);
  if (rc) return rc;
  __nv_bfloat16* ut = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(ws) + WS_CTR_BYTES + WS_PROF_BYTES);
  for (int d = 0; d < ND; ++d) {
    rc = ds2_transpose(static_cast<const __nv_bfloat16*>(u) + (size_t)d * H * H, DS2_BF16, H, H, H, ut + (size_t)d * H * H, DS2_BF16, H, st);
    if (rc) return rc;
  }
  CUtensorMap ta;
  rc = make_tmap_2d_bf16(&ta, ut, H, (uint64_t)ND * H, H, 64, 128);
  if (rc) return rc;
  p.add_in = dout; p.bias = nullptr; p.lens = lens; p.hs_in = static_cast<const __nv_bfloat16*>(hs);
  p.dbias = dbias;
  if (dbias) DS2_CUDA(cudaMemsetAsync(dbias, 0, (size_t)ND * H * sizeof(float), st));
  p.out = static_cast<__nv_bfloat16*>(dz);
  p.out_row_pitch = (long)ND * H; p.out_dir_stride = H;
  p.T = T; p.B = B; p.H = H; p.ND = ND; p.reverse_mask = reverse_mask; p.scale = 1.f; p.cap = cap;
  if (!p.fenced) DS2_CUDA(cudaMemsetAsync(dz, DS2_FILL_BYTE, (size_t)T * B * ND * H * 2, st));
  return persist_launch<true>(ta, p, smem, st);
}

}  // namespace ds2
