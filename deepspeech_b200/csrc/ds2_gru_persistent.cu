// Persistent, weights-resident GRU recurrence (gate order r,z,n; cuDNN / torch.nn.GRU form), forward and backward.
// GRU is NOT in the reference: BASELINE configs 3 and 5 ask for it and the build defines it
// (include/ds2_b200.h).  Same machine as ds2_rnn_persistent.cu:
//   * hidden units tiled by 128 (tcgen05 M), a 16-utterance instance is the MMA N, K split over a 4-CTA cluster;
//   * forward: the CTA keeps THREE [128 x H/4] bf16 slices of U (gates r, z, n of its 128 units) resident in TMEM
//     and produces three accumulator tiles per step; the gate math of a unit needs all three, so a CTA finishes
//     the same 32 units of every gate:  r,z = sigmoid(s*gi + bw + U h + bu), n = tanh(s*gi_n + bw_n + r*(U_n h + bu_n)),
//     h' = (1-z) n + z h;  (r, z, n, U_n h + bu_n) are stored in bf16 for the backward;
//   * backward: A = U^T [H x 3H] (K = 3H split over the cluster), moving operand dgh_{t+1};
//     dh = dout_t + z_{t+1}-carry + U^T dgh_{t+1};  dn = dh (1-z)(1-n^2), dz = dh (h_prev - n) z (1-z), dr = dn ghn r (1-r);
//     dgi = (dr, dz, dn) feeds dW / dx / dbw, dgh = (dr, dz, dn r) feeds dU / dbu and the next step;
//   * step ordering, operand staging, slab exchange: see ds2_rnn_persistent.cu.
#include <stdlib.h>

#include "common.cuh"

namespace ds2 {

// cluster size = K split: 4 when the CTA's K slices of U fit its 512 TMEM columns, else 8 (H = 1760: the forward keeps
// 3 gates x 256 of 1760 columns, the backward 704 of 5280; a CTA then finishes 16 units instead of 32)
static constexpr int GBP = 16;              // utterances per instance (MMA N)
static constexpr int GINST_THREADS = 160;   // per instance: 1 control warp + 4 loader/epilogue warps
static constexpr int GMAX_NI = 2;
static constexpr int GMAX_NBOX = 12;        // K-slice <= 768 columns
static constexpr int GSTAGE_BOX = 4;        // boxes of U staged through shared memory at a time (64 KB)
static constexpr int GCHUNK = GBP * GMAX_NI; // utterances per batch chunk (one cluster column of CTAs each)

struct GruParams {
  const float* add_in;          // fwd: wx [T,B,ND*3H] ; bwd: dout [T,B,H]
  const float* bw;              // fwd: [ND,3H]
  const float* bu;              // fwd: [ND,3H]
  const int* lens;              // [B]
  __nv_bfloat16* hs;            // [ND,T,B,H]   fwd: out (and moving operand) ; bwd: in
  __nv_bfloat16* gates;         // [ND,T,B,4H]  fwd: out ; bwd: in   (r | z | n | U_n h + bu_n)
  __nv_bfloat16* dgi;           // bwd out [T,B,ND*3H]
  __nv_bfloat16* dgh;           // bwd out [T,B,ND*3H] (and moving operand)
  float* dbw;                   // bwd, optional: [ND,3H] receives sum_{t,b} dgi (zero-filled by the host side)
  float* dbu;                   // bwd, optional: [ND,3H] receives sum_{t,b} dgh
  unsigned int* ctr;            // [NC][ND][NI] step counters, 32 uints apart
  long long* prof;              // optional (DS2_RNN_PROF=1): 8 clock stamps per step of CTA (0,0,0), instance 0, first loader thread
  int T, B, H, ND, reverse_mask;
  int NC, NI, tiles_m, ksl, nbox;   // batch chunks, instances per chunk, M tiles per direction, K-slice length
                                    // (multiple of 64), boxes per slice
  int z_base;                       // first (chunk, direction) pair of this launch: pair z = chunk * ND + d
  float scale;
  int fenced, slab16;           // see RnnPersistParams (ds2_rnn_persistent.cu)
  int bulk;                     // forward: a warp parks its bf16 partial sums in smem and sends them with one bulk copy
};

__device__ __forceinline__ uint32_t g_cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void g_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t g_map_to_rank(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void g_st_async_v4(uint32_t remote_addr, uint32_t remote_bar, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%2, %3, %4, %5}, [%1];"
               ::"r"(remote_addr), "r"(remote_bar), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 g_ld_relaxed_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int g_ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void g_red_release_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void g_tmem_cp_128x256b(uint32_t taddr, uint64_t sdesc) {
  asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void g_umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void g_tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ float g_f4(const float4& v, int i) { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }
__device__ __forceinline__ float g_sigmoid(float x) { return 1.f / (1.f + __expf(-x)); }
__device__ __forceinline__ void g_unpack4(uint2 v, float (&o)[4]) {
  const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) o[i] = __bfloat162float(h[i]);
}
__device__ __forceinline__ uint2 g_pack4(const float (&o)[4]) {
  uint2 pk;
  __nv_bfloat162 a = __floats2bfloat162_rn(o[0], o[1]), b = __floats2bfloat162_rn(o[2], o[3]);
  pk.x = scrub_fill(*reinterpret_cast<uint32_t*>(&a));
  pk.y = scrub_fill(*reinterpret_cast<uint32_t*>(&b));
  return pk;
}

struct GruSmem {
  uint32_t a_off, b_off, slab_off, bar_off, total;
};
static constexpr int GBARS_PER_INST = 3;   // b_ready, mma_done, slab_full
__host__ __device__ inline GruSmem gru_smem_layout(int nbox, int NI, int NG) {   // the slab size does not depend on the K split
  GruSmem s;
  s.a_off = 0;                                              // GSTAGE_BOX x [128 x 64] bf16 staging of U
  s.b_off = s.a_off + GSTAGE_BOX * 16384;                         // NI x nbox x [16 x 64] bf16
  s.slab_off = s.b_off + NI * nbox * (GBP * 128);           // NI x KS x [32 x NG*16] fp32
  s.bar_off = s.slab_off + NI * 128 * NG * GBP * 4;
  s.total = s.bar_off + 8 * (2 + NI * GBARS_PER_INST) + 16;
  return s;
}

// NG accumulator / A-operand groups: 3 gates (forward) or 1 (backward)
template <bool BWD, int KSZ>
__global__ void __launch_bounds__(GINST_THREADS * GMAX_NI, 1)
gru_persist_kernel(const __grid_constant__ CUtensorMap tm_a, GruParams p) {
  constexpr int NG = BWD ? 1 : 3;
  constexpr int UPC = 128 / KSZ;            // hidden units this CTA finishes per tile
  // 16-byte units of the moving operand per loader thread (= nbox): the forward's three resident gate slices
  // cap nbox at 4 (TMEM columns), the backward's single slice of the 3H-long contraction at GMAX_NBOX
  constexpr int GMAX_UNITS = BWD ? GMAX_NBOX : 4;
  extern __shared__ uint8_t gru_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(gru_smem_raw) + 1023) & ~uintptr_t(1023));
  const int NI = p.NI, nbox = p.nbox;
  const GruSmem L = gru_smem_layout(nbox, NI, NG);
  uint8_t* sA = smem + L.a_off;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
  uint64_t* a_bar = bars;            // a gate's U slice landed in smem (one phase per gate)
  uint64_t* acp_bar = bars + 1;      // ... and was copied into TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 + NI * GBARS_PER_INST);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int inst = warp / 5, w_in = warp % 5;
  uint64_t* b_ready = bars + 2 + inst * GBARS_PER_INST;
  uint64_t* mma_bar = b_ready + 1;
  uint64_t* slab_bar = b_ready + 2;
  uint8_t* sB = smem + L.b_off + inst * nbox * (GBP * 128);
  float* slab = reinterpret_cast<float*>(smem + L.slab_off) + inst * 128 * NG * GBP;
  // shared-space addresses used inside the step loop, computed once
  const uint32_t b_ready_a = keep_reg(smem_u32(b_ready)), mma_bar_a = b_ready_a + 8, slab_bar_a = b_ready_a + 16;
  const uint32_t sB_addr = keep_reg(smem_u32(sB)), slab_local = keep_reg(smem_u32(slab));

  const uint32_t rank = g_cluster_ctarank();
  const int zz = p.z_base + blockIdx.z;
  const int mt = blockIdx.y, chunk = zz / p.ND, d = zz % p.ND;
  const int m0 = mt * 128;
  const int rev = (p.reverse_mask >> d) & 1;
  const int T = p.T, H = p.H, ND = p.ND, Bn = p.B;
  const int KT = BWD ? 3 * H : H;                           // length of the contraction
  unsigned int* ctr = p.ctr + (blockIdx.z * NI + inst) * 32;
  const unsigned int ctas_per_dir = KSZ * p.tiles_m;
  const int acols = NG * nbox * 32;                         // TMEM columns holding U

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm_a);
    mbar_init(a_bar, 1);
    mbar_init(acp_bar, 1);
    for (int i = 0; i < NI; ++i) {
      uint64_t* bb = bars + 2 + i * GBARS_PER_INST;
      mbar_init(bb, 128);
      mbar_init(bb + 1, 1);
      mbar_init(bb + 2, 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t d_tmem = tmem_base + acols + inst * NG * GBP;
  g_cluster_sync();

  // resident-operand set-up: the CTA's NG x nbox boxes [128 x 64] of U go through a GSTAGE_BOX-deep staging
  // buffer into TMEM (box x of gate g -> columns (g*nbox + x)*32 ...)
  if (warp == 0) {
    if (elect_one_sync()) {
      const uint32_t a0 = smem_u32(sA);
      const int total = NG * nbox;
      uint32_t sph = 0;
      for (int c0 = 0; c0 < total; c0 += GSTAGE_BOX) {
        const int nb = min(GSTAGE_BOX, total - c0);
        mbar_arrive_expect_tx(a_bar, nb * 16384);
        for (int x = 0; x < nb; ++x) {
          const int g = (c0 + x) / nbox, i = (c0 + x) % nbox;
          // forward: gate g of this tile's 128 units = rows d*3H + g*H + m0 of U;  backward: rows d*H + m0 of U^T
          const int row0 = BWD ? d * H + m0 : d * 3 * H + g * H + m0;
          tma_load_2d(sA + x * 16384, &tm_a, a_bar, rank * p.ksl + i * 64, row0);
        }
        mbar_wait(a_bar, sph, 10);
        tc_fence_after();
        for (int c = 0; c < nb * 4; ++c)
          g_tmem_cp_128x256b(tmem_base + (c0 * 4 + c) * 8, umma_smem_desc_sw128(a0 + (c >> 2) * 16384 + (c & 3) * 32, 16, 1024));
        umma_commit(acp_bar);
        mbar_wait(acp_bar, sph, 11);        // copies done: the staging buffer is free again
        sph ^= 1;
      }
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (w_in == 0) {
    // =========================== control warp of this instance: MMA issue ===========================
    const uint32_t idesc = umma_idesc_bf16(128, GBP, 0, 0);
    const uint32_t b0 = sB_addr;
    uint32_t ph = 0;
    for (int step = 1; step < T; ++step) {
      mbar_wait(b_ready_a, ph, 30);
      tc_fence_after();
      if (elect_one_sync()) {
#pragma unroll
        for (int g = 0; g < NG; ++g) {
          for (int i = 0; i < nbox; ++i) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              g_umma_bf16_ts(d_tmem + g * GBP, tmem_base + ((g * nbox + i) * 4 + k) * 8,
                             umma_smem_desc_sw128(b0 + i * (GBP * 128) + k * 32, 16, 1024), idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(mma_bar_a);
      }
      __syncwarp();
      ph ^= 1;
    }
  } else {
    // =========================== loader + epilogue warps ===========================
    const int q = warp & 3;
    const int et = (w_in - 1) * 32 + lane;       // 0..127
    const int bl = et % GBP;
    const int jg = et / GBP;                     // 0..7
    const int b = chunk * GCHUNK + inst * GBP + bl;
    const int j = m0 + UPC * (int)rank + 4 * jg; // first of 4 consecutive hidden units
    const bool jvalid = jg < UPC / 4 && j < H;
    const int len_b = p.lens[b];
    const uint32_t slab_bar_local = slab_bar_a;
    // moving operand: forward h_{t-1} (rows of hs[d]), backward dgh_{t+1} (rows of dgh, direction offset d*3H)
    const __nv_bfloat16* src_dir = BWD ? p.dgh + (size_t)d * 3 * H : p.hs + (size_t)d * T * Bn * H;
    const long src_pitch = BWD ? (long)ND * 3 * H : (long)H;
    const int cpr = p.ksl / 8;
    const int nunits = GBP * cpr;
    int goff[GMAX_UNITS];
    int soff[GMAX_UNITS];
#pragma unroll
    for (int m = 0; m < GMAX_UNITS; ++m) {
      const int u = et + 128 * m;
      goff[m] = -1;
      soff[m] = -1;
      if (u < nunits) {
        const int ub = u / cpr, uc = u - ub * cpr;
        soff[m] = (uc >> 3) * (GBP * 128) + ub * 128 + (((uc & 7) ^ (ub & 7)) << 4);
        if ((int)rank * p.ksl + uc * 8 < KT) goff[m] = ub * (int)src_pitch + (int)rank * p.ksl + uc * 8;
      }
    }
    // biases of this thread's units: bw+bu folded for the r and z gates; n keeps them apart (r multiplies U_n h + bu_n)
    float cb[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const bool ok = !BWD && jvalid;
      const size_t o = (size_t)d * 3 * H + j + i;
      cb[0][i] = ok ? p.bw[o] + p.bu[o] : 0.f;
      cb[1][i] = ok ? p.bw[o + H] + p.bu[o + H] : 0.f;
      cb[2][i] = ok ? p.bw[o + 2 * (size_t)H] : 0.f;
      cb[3][i] = ok ? p.bu[o + 2 * (size_t)H] : 0.f;
    }
    // forward: h_{t-1} of this thread's (b, units); backward: dh_{t+1} * z_{t+1}, the direct path of dh
    float carry[4] = {0.f, 0.f, 0.f, 0.f};
    float sdr[4] = {0.f, 0.f, 0.f, 0.f}, sdz[4] = {0.f, 0.f, 0.f, 0.f}, sdn[4] = {0.f, 0.f, 0.f, 0.f}, sdnr[4] = {0.f, 0.f, 0.f, 0.f};
    const bool pr = p.prof != nullptr && blockIdx.x + blockIdx.y + blockIdx.z == 0 && inst == 0 && et == 0;
    uint32_t ph = 0;
    for (int step = 0; step < T; ++step) {
      if (pr && step < 1024) p.prof[step * 8 + 0] = dev_clock();
      const int t = BWD ? (rev ? step : T - 1 - step) : (rev ? T - 1 - step : step);
      const int tsrc = BWD ? (rev ? t - 1 : t + 1) : (rev ? t + 1 : t - 1);      // producer of the moving operand
      const int tprev = rev ? t + 1 : t - 1;                                     // previous step in FORWARD order
      const bool has_prev = tprev >= 0 && tprev < T;
      // ---- prefetch this step's additive inputs
      float4 gi[3];
      float4 dout4 = make_float4(0.f, 0.f, 0.f, 0.f);
      uint2 g4[4];
      uint2 hp2 = make_uint2(0u, 0u);             // backward: h of the previous step in forward order (zero before the first)
      if (jvalid) {
        if (!BWD) {
          const float* src = p.add_in + ((size_t)t * Bn + b) * ND * 3 * H + (size_t)d * 3 * H + j;
#pragma unroll
          for (int g = 0; g < 3; ++g) gi[g] = __ldg(reinterpret_cast<const float4*>(src + (size_t)g * H));
        } else {
          dout4 = __ldg(reinterpret_cast<const float4*>(p.add_in + ((size_t)t * Bn + b) * H + j));
          const __nv_bfloat16* gp = p.gates + (((size_t)d * T + t) * Bn + b) * 4 * H + j;
#pragma unroll
          for (int g = 0; g < 4; ++g) g4[g] = *reinterpret_cast<const uint2*>(gp + (size_t)g * H);
          if (has_prev) hp2 = *reinterpret_cast<const uint2*>(p.hs + (((size_t)d * T + tprev) * Bn + b) * H + j);
        }
      }
      float acc[NG][4];
#pragma unroll
      for (int g = 0; g < NG; ++g)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[g][i] = 0.f;
      if (step > 0) {
        if (et == 0) mbar_arrive_expect_tx(slab_bar_a, p.slab16 ? 128 * NG * GBP * 2 : 128 * NG * GBP * 4);
        if (et == 0) {
          const unsigned int need = (p.fenced ? 1u : 4u) * ctas_per_dir * (unsigned int)step;   // unfenced: every epilogue warp publishes
          if (p.fenced) {
            if (g_ld_acquire_u32(ctr) < need) {
              SpinGuard sg;
              while (g_ld_acquire_u32(ctr) < need) sg.check(20);
            }
          } else if (ld_relaxed_gpu_u32(ctr) < need) {
            SpinGuard sg;
            while (ld_relaxed_gpu_u32(ctr) < need) sg.check(21);
          }
        }
        asm volatile("bar.sync %0, 128;" ::"r"(1 + inst) : "memory");
        if (pr && step < 1024) p.prof[step * 8 + 1] = dev_clock();
        {
          const __nv_bfloat16* rows = src_dir + ((size_t)tsrc * Bn + chunk * GCHUNK + inst * GBP) * src_pitch;
          uint4 v[GMAX_UNITS];
#pragma unroll
          for (int m = 0; m < GMAX_UNITS; ++m) {
            v[m] = make_uint4(0u, 0u, 0u, 0u);
            if (goff[m] >= 0) v[m] = g_ld_relaxed_v4(rows + goff[m]);
          }
          if (!p.fenced) {
            // relaxed step counter: a row may still show the host's fill (common.cuh); reload until it landed
#pragma unroll
            for (int m = 0; m < GMAX_UNITS; ++m) {
              if (goff[m] >= 0 && unit_stale(v[m])) {
                SpinGuard sg;
                do {
                  v[m] = g_ld_relaxed_v4(rows + goff[m]);
                  sg.check(22);
                } while (unit_stale(v[m]));
              }
            }
          }
#pragma unroll
          for (int m = 0; m < GMAX_UNITS; ++m)
            if (soff[m] >= 0) sts_v4(sB_addr + soff[m], v[m]);
          fence_proxy_async();
          mbar_arrive(b_ready_a);
        }
        if (pr && step < 1024) p.prof[step * 8 + 2] = dev_clock();
        mbar_wait(mma_bar_a, ph, 40);
        tc_fence_after();
        if (pr && step < 1024) p.prof[step * 8 + 3] = dev_clock();
        // partial sums of this CTA's K slice -> the cluster CTA that finishes those units (all gates)
        if (!BWD && p.slab16 && p.bulk) {
          // Three gates are 6 st.async per thread: the DSMEM store queue backs up into the LSU (1700 clk to issue, and the
          // epilogue's own loads and stores queue behind it).  Instead the warp writes its 32 rows x 3 gates x 16 bf16
          // (3 KB, already in the slab's layout) to local shared memory and one thread sends them with a bulk copy per
          // finishing CTA, completing on that CTA's slab barrier like the st.async did.  The staging area is the U staging
          // buffer, free after set-up; a step's copy has been consumed before any CTA can reach the next step.
          const uint32_t stg = smem_u32(sA) + (inst * 4 + (w_in - 1)) * (32 * NG * GBP * 2);
#pragma unroll
          for (int g = 0; g < NG; ++g) {
            uint32_t r[16];
            g_tmem_ld_32x16(d_tmem + g * GBP + (static_cast<uint32_t>(q * 32) << 16), r);
            tmem_ld_wait();
            uint32_t h[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(r[2 * c]), __uint_as_float(r[2 * c + 1]));
              h[c] = *reinterpret_cast<uint32_t*>(&t2);
            }
            sts_v4(stg + (lane * NG + g) * (GBP * 2), make_uint4(h[0], h[1], h[2], h[3]));
            sts_v4(stg + (lane * NG + g) * (GBP * 2) + 16, make_uint4(h[4], h[5], h[6], h[7]));
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
#pragma unroll
            for (int hh = 0; hh < 32 / UPC; ++hh) {
              const uint32_t to = (q * 32 + hh * UPC) / UPC;
              const uint32_t dstb = g_map_to_rank(slab_local, to) + rank * (UPC * NG * GBP * 2);
              const uint32_t rbar = g_map_to_rank(slab_bar_local, to);
              asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                           ::"r"(dstb), "r"(stg + hh * (UPC * NG * GBP * 2)), "r"(UPC * NG * GBP * 2), "r"(rbar) : "memory");
            }
          }
        } else
#pragma unroll
        for (int g = 0; g < NG; ++g) {
          uint32_t r[16];
          g_tmem_ld_32x16(d_tmem + g * GBP + (static_cast<uint32_t>(q * 32) << 16), r);
          tmem_ld_wait();
          const uint32_t row = q * 32 + lane, to = row / UPC;       // tile row -> finishing CTA
          const uint32_t off = (((rank * UPC + row % UPC) * NG + g) * GBP) * 4;
          const uint32_t dst = g_map_to_rank(slab_local, to) + off;
          const uint32_t rbar = g_map_to_rank(slab_bar_local, to);
          if (p.slab16) {
            uint32_t h[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(r[2 * c]), __uint_as_float(r[2 * c + 1]));
              h[c] = *reinterpret_cast<uint32_t*>(&t2);
            }
            const uint32_t dst2 = g_map_to_rank(slab_local, to) + off / 2;
            g_st_async_v4(dst2, rbar, h[0], h[1], h[2], h[3]);
            g_st_async_v4(dst2 + 16, rbar, h[4], h[5], h[6], h[7]);
          } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) g_st_async_v4(dst + c * 16, rbar, r[4 * c], r[4 * c + 1], r[4 * c + 2], r[4 * c + 3]);
          }
        }
        tc_fence_before();
        if (pr && step < 1024) p.prof[step * 8 + 4] = dev_clock();
        mbar_wait(slab_bar_a, ph, 50);
        if (pr && step < 1024) p.prof[step * 8 + 5] = dev_clock();
#pragma unroll
        for (int s = 0; s < KSZ; ++s)
#pragma unroll
          for (int g = 0; g < NG; ++g)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              if (jg < UPC / 4) {
                const uint32_t e = ((s * UPC + 4 * jg + i) * NG + g) * GBP + bl;
                if (p.slab16) {
                  unsigned short hv;
                  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(hv) : "r"(slab_local + e * 2) : "memory");
                  acc[g][i] += __uint_as_float((uint32_t)hv << 16);
                } else {
                  acc[g][i] += lds_f32(slab_local + e * 4);
                }
              }
            }
        ph ^= 1;
      }
      uint2 late[4];
      if (jvalid) {
        const bool live = t < len_b;
        // h of the previous step in forward order (zero before the first step)
        // forward: this thread produced h_prev of its own (b, units) last step and still holds it (bf16-rounded,
        // as stored); backward: read it back from hs
        float hp[4] = {0.f, 0.f, 0.f, 0.f};
        if (!BWD) {
#pragma unroll
          for (int i = 0; i < 4; ++i) hp[i] = carry[i];
        } else {
          g_unpack4(hp2, hp);                     // fetched with the other inputs at the top of the step
        }
        if (!BWD) {
          float ho[4], ro[4], zo[4], no[4], go[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float r = g_sigmoid(fmaf(p.scale, g_f4(gi[0], i), cb[0][i]) + acc[0][i]);
            const float z = g_sigmoid(fmaf(p.scale, g_f4(gi[1], i), cb[1][i]) + acc[1][i]);
            const float ghn = acc[2][i] + cb[3][i];
            const float n = tanhf(fmaf(p.scale, g_f4(gi[2], i), cb[2][i]) + r * ghn);
            ho[i] = live ? (1.f - z) * n + z * hp[i] : 0.f;
            ro[i] = r; zo[i] = z; no[i] = n; go[i] = ghn;
          }
          const uint2 hpk = g_pack4(ho);
          g_unpack4(hpk, carry);
          *reinterpret_cast<uint2*>(p.hs + (((size_t)d * T + t) * Bn + b) * H + j) = hpk;
          // (r, z, n, U_n h + b) are only read by the backward pass: stored after the publication
          late[0] = g_pack4(ro); late[1] = g_pack4(zo); late[2] = g_pack4(no); late[3] = g_pack4(go);
        } else {
          float r4[4], z4[4], n4[4], gh4[4];
          g_unpack4(g4[0], r4); g_unpack4(g4[1], z4); g_unpack4(g4[2], n4); g_unpack4(g4[3], gh4);
          const float dv[4] = {dout4.x, dout4.y, dout4.z, dout4.w};
          float dr[4], dz[4], dn[4], dnr[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float dh = live ? dv[i] + carry[i] + acc[0][i] : 0.f;
            dn[i] = dh * (1.f - z4[i]) * (1.f - n4[i] * n4[i]);
            dz[i] = dh * (hp[i] - n4[i]) * z4[i] * (1.f - z4[i]);
            dr[i] = dn[i] * gh4[i] * r4[i] * (1.f - r4[i]);
            dnr[i] = dn[i] * r4[i];
            carry[i] = dh * z4[i];
            sdr[i] += dr[i]; sdz[i] += dz[i]; sdn[i] += dn[i]; sdnr[i] += dnr[i];
          }
          __nv_bfloat16* gip = p.dgi + ((size_t)t * Bn + b) * ND * 3 * H + (size_t)d * 3 * H + j;
          __nv_bfloat16* ghp = p.dgh + ((size_t)t * Bn + b) * ND * 3 * H + (size_t)d * 3 * H + j;
          // dgh is what the other CTAs wait for; dgi (read by the GEMMs after the kernel) follows the publication
          late[0] = g_pack4(dr); late[1] = g_pack4(dz); late[2] = g_pack4(dn);
          *reinterpret_cast<uint2*>(ghp) = late[0];
          *reinterpret_cast<uint2*>(ghp + (size_t)H) = late[1];
          *reinterpret_cast<uint2*>(ghp + (size_t)2 * H) = g_pack4(dnr);
          (void)gip;
        }
      }
      if (pr && step < 1024) p.prof[step * 8 + 6] = dev_clock();
      if (p.fenced) {
        asm volatile("bar.sync %0, 128;" ::"r"(1 + inst) : "memory");
        if (et == 0) g_red_release_add(ctr, 1u);
      } else {
        // no CTA barrier: every warp publishes its own rows (the data validates itself, common.cuh)
        __syncwarp();
        if (lane == 0) red_relaxed_gpu_add(ctr, 1u);
      }
      if (jvalid) {
        if (!BWD) {
          __nv_bfloat16* gp = p.gates + (((size_t)d * T + t) * Bn + b) * 4 * H + j;
          *reinterpret_cast<uint2*>(gp) = late[0];
          *reinterpret_cast<uint2*>(gp + (size_t)H) = late[1];
          *reinterpret_cast<uint2*>(gp + (size_t)2 * H) = late[2];
          *reinterpret_cast<uint2*>(gp + (size_t)3 * H) = late[3];
        } else {
          __nv_bfloat16* gip = p.dgi + ((size_t)t * Bn + b) * ND * 3 * H + (size_t)d * 3 * H + j;
          *reinterpret_cast<uint2*>(gip) = late[0];
          *reinterpret_cast<uint2*>(gip + (size_t)H) = late[1];
          *reinterpret_cast<uint2*>(gip + (size_t)2 * H) = late[2];
        }
      }
      if (pr && step < 1024) p.prof[step * 8 + 7] = dev_clock();
    }
    if (BWD && p.dbw != nullptr) {
      // bias gradients: sum over the instance's 16 utterances (16 consecutive lanes share a unit quad), one red per unit
      // dbw = sum (dr, dz, dn), dbu = sum (dr, dz, dn*r)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
          sdr[i] += __shfl_xor_sync(0xffffffffu, sdr[i], o);
          sdz[i] += __shfl_xor_sync(0xffffffffu, sdz[i], o);
          sdn[i] += __shfl_xor_sync(0xffffffffu, sdn[i], o);
          sdnr[i] += __shfl_xor_sync(0xffffffffu, sdnr[i], o);
        }
      }
      if (bl == 0 && jvalid) {
        float* w = p.dbw + (size_t)d * 3 * H + j;
        float* u = p.dbu + (size_t)d * 3 * H + j;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          atomicAdd(w + i, sdr[i]); atomicAdd(w + H + i, sdz[i]); atomicAdd(w + 2 * (size_t)H + i, sdn[i]);
          atomicAdd(u + i, sdr[i]); atomicAdd(u + H + i, sdz[i]); atomicAdd(u + 2 * (size_t)H + i, sdnr[i]);
        }
      }
    }
  }
  tc_fence_before();
  g_cluster_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------ host side
static constexpr size_t GWS_CTR_BYTES = 2048;
static constexpr size_t GWS_PROF_BYTES = 65536;     // 1024 steps x 8 stamps
struct GruGeom {
  int ks;              // cluster size (K split): 4 or 8
  int tiles_m, ksl, nbox, ni, nc;
  int zper;            // (chunk, direction) pairs one launch can hold: all of its CTAs must be resident at once
  size_t smem;
};
static int gru_geometry(bool bwd, int B, int H, int ND, GruGeom* g) {
  // whole 16-utterance instances; more than one chunk of 32 only in whole chunks
  if (B < GBP || B % GBP != 0 || (B > GCHUNK && B % GCHUNK != 0) || H % 8 != 0 || H < 8 || (ND != 1 && ND != 2)) return 0;
  const int NG = bwd ? 1 : 3;
  const int KT = bwd ? 3 * H : H;
  const int tm = (H + 127) / 128;
  const int n_chunks = B <= GCHUNK ? 1 : B / GCHUNK;
  const int n_inst = B <= GCHUNK ? B / GBP : GMAX_NI;
  for (int ks = 4; ks <= 8; ks *= 2) {
    const int slice = (KT + ks - 1) / ks;
    const int l = ((slice + 63) / 64) * 64;
    const int nb = l / 64;
    if (nb > (bwd ? GMAX_NBOX : 4)) continue;
    if (NG * nb * 32 + n_inst * NG * GBP > 512) continue;      // TMEM columns: resident U + accumulators
    const size_t bytes = gru_smem_layout(nb, n_inst, NG).total + 1024;
    if (bytes > 225 * 1024) continue;
    // every CTA of a launch must be resident at once (they wait on each other): one CTA per SM, 148 SMs
    int zper = 144 / (ks * tm);
    if (zper < 1) continue;
    if (zper > ND * n_chunks) zper = ND * n_chunks;
    if (zper * n_inst * 32 * 4 > (int)GWS_CTR_BYTES) zper = (int)GWS_CTR_BYTES / (n_inst * 32 * 4);
    if (g) {
      g->ks = ks; g->tiles_m = tm; g->ksl = l; g->nbox = nb; g->ni = n_inst; g->nc = n_chunks; g->zper = zper; g->smem = bytes;
    }
    return 1;
  }
  return 0;
}

int gru_persist_supported(int B, int H, int ND) {
  return gru_geometry(false, B, H, ND, nullptr) && gru_geometry(true, B, H, ND, nullptr);
}

size_t gru_persist_workspace_bytes(int T, int B, int H, int ND) {
  (void)T; (void)B;
  return GWS_CTR_BYTES + GWS_PROF_BYTES + (size_t)ND * 3 * H * H * 2;          // counters + stamps + U^T (bf16) for the backward
}

template <bool BWD, int KSZ>
static int gru_launch_ks(const CUtensorMap& ta, GruParams& p, const GruGeom& g, cudaStream_t st) {
  auto kern = gru_persist_kernel<BWD, KSZ>;
  static unsigned long long attr_done_devs = 0;
  if (first_use_on_device(&attr_done_devs)) {
    DS2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  }
  // one launch per group of (chunk, direction) pairs that fits the device; groups run one after the other
  const int total = p.ND * p.NC;
  for (int z0 = 0; z0 < total; z0 += g.zper) {
    DS2_CUDA(cudaMemsetAsync(p.ctr, 0, GWS_CTR_BYTES, st));
    p.z_base = z0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(KSZ, p.tiles_m, total - z0 < g.zper ? total - z0 : g.zper);
    cfg.blockDim = dim3(GINST_THREADS * p.NI);
    cfg.dynamicSmemBytes = g.smem;
    cfg.stream = st;
    cudaLaunchAttribute attrs[2];
    attrs[0].id = cudaLaunchAttributeClusterDimension;
    attrs[0].val.clusterDim.x = KSZ;
    attrs[0].val.clusterDim.y = 1;
    attrs[0].val.clusterDim.z = 1;
    attrs[1].id = cudaLaunchAttributeCooperative;
    attrs[1].val.cooperative = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = use_cooperative_launch() ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, p);
    if (e != cudaSuccess && cfg.numAttrs == 2) {
      (void)cudaGetLastError();
      cfg.numAttrs = 1;
      e = cudaLaunchKernelEx(&cfg, kern, ta, p);
    }
    if (e != cudaSuccess) return cuda_fail(e, BWD ? "gru_persist_kernel<bwd>" : "gru_persist_kernel<fwd>");
    count_launch();
  }
  return DS2_OK;
}

template <bool BWD>
static int gru_launch(const CUtensorMap& ta, GruParams& p, const GruGeom& g, cudaStream_t st) {
  p.tiles_m = g.tiles_m; p.ksl = g.ksl; p.nbox = g.nbox; p.NI = g.ni; p.NC = g.nc;
  return g.ks == 4 ? gru_launch_ks<BWD, 4>(ta, p, g, st) : gru_launch_ks<BWD, 8>(ta, p, g, st);
}

int gru_fwd_persist(const float* wx, float scale, const void* u, const float* bw, const float* bu, const int* lens, void* hs,
                    void* gates, int T, int B, int H, int ND, int reverse_mask, void* ws, size_t ws_bytes, cudaStream_t st) {
  GruParams p = {};
  GruGeom geo;
  DS2_REQUIRE(gru_geometry(false, B, H, ND, &geo), DS2_ERR_UNSUPPORTED,
              "ds2_gru_fwd: persistent kernel does not take B=%d H=%d ND=%d", B, H, ND);
  DS2_REQUIRE(ws && ws_bytes >= GWS_CTR_BYTES + GWS_PROF_BYTES, DS2_ERR_WORKSPACE, "ds2_gru_fwd: workspace of %zu bytes required",
              GWS_CTR_BYTES + GWS_PROF_BYTES);
  p.prof = env().rnn_prof ? reinterpret_cast<long long*>(static_cast<char*>(ws) + GWS_CTR_BYTES) : nullptr;
  CUtensorMap ta;
  int rc = make_tmap_2d_bf16(&ta, u, H, (uint64_t)ND * 3 * H, H, 64, 128);
  if (rc) return rc;
  p.add_in = wx; p.bw = bw; p.bu = bu; p.lens = lens; p.hs = static_cast<__nv_bfloat16*>(hs);
  p.gates = static_cast<__nv_bfloat16*>(gates); p.dgi = nullptr; p.dgh = nullptr;
  p.ctr = static_cast<unsigned int*>(ws);
  p.T = T; p.B = B; p.H = H; p.ND = ND; p.reverse_mask = reverse_mask; p.scale = scale;
  p.fenced = env().rnn_fenced;
  p.slab16 = env().rnn_slab_f32 ? 0 : 1;
  p.bulk = env().gru_no_bulk ? 0 : 1;
  if (!p.fenced) DS2_CUDA(cudaMemsetAsync(hs, DS2_FILL_BYTE, (size_t)ND * T * B * H * 2, st));
  return gru_launch<false>(ta, p, geo, st);
}

int gru_bwd_persist(const float* dout, const void* u, const int* lens, const void* hs, const void* gates, void* dgi, void* dgh,
                    float* dbw, float* dbu, int T, int B, int H, int ND, int reverse_mask, void* ws, size_t ws_bytes, cudaStream_t st) {
  GruParams p = {};
  GruGeom geo;
  DS2_REQUIRE(gru_geometry(true, B, H, ND, &geo), DS2_ERR_UNSUPPORTED,
              "ds2_gru_bwd: persistent kernel does not take B=%d H=%d ND=%d", B, H, ND);
  const size_t need = gru_persist_workspace_bytes(T, B, H, ND);
  DS2_REQUIRE(ws && ws_bytes >= need, DS2_ERR_WORKSPACE, "ds2_gru_bwd: workspace of %zu bytes required", need);
  // U^T per direction: ut[d][i][g*H + j] = U[d][g*H + j][i]
  __nv_bfloat16* ut = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(ws) + GWS_CTR_BYTES + GWS_PROF_BYTES);
  p.prof = env().rnn_prof ? reinterpret_cast<long long*>(static_cast<char*>(ws) + GWS_CTR_BYTES) : nullptr;
  int rc;
  for (int d = 0; d < ND; ++d) {
    rc = ds2_transpose(static_cast<const __nv_bfloat16*>(u) + (size_t)d * 3 * H * H, DS2_BF16, 3L * H, H, H,
                       ut + (size_t)d * 3 * H * H, DS2_BF16, 3L * H, st);
    if (rc) return rc;
  }
  CUtensorMap ta;
  rc = make_tmap_2d_bf16(&ta, ut, 3L * H, (uint64_t)ND * H, 3L * H, 64, 128);
  if (rc) return rc;
  p.add_in = dout; p.bw = nullptr; p.bu = nullptr; p.lens = lens;
  p.hs = const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(hs));
  p.gates = const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(gates));
  p.dgi = static_cast<__nv_bfloat16*>(dgi); p.dgh = static_cast<__nv_bfloat16*>(dgh);
  p.dbw = (dbw && dbu) ? dbw : nullptr; p.dbu = dbu;
  if (p.dbw) {
    DS2_CUDA(cudaMemsetAsync(dbw, 0, (size_t)ND * 3 * H * sizeof(float), st));
    DS2_CUDA(cudaMemsetAsync(dbu, 0, (size_t)ND * 3 * H * sizeof(float), st));
  }
  p.ctr = static_cast<unsigned int*>(ws);
  p.T = T; p.B = B; p.H = H; p.ND = ND; p.reverse_mask = reverse_mask; p.scale = 1.f;
  p.fenced = env().rnn_fenced;
  p.slab16 = env().rnn_slab_f32 ? 0 : 1;
  p.bulk = env().gru_no_bulk ? 0 : 1;
  if (!p.fenced) DS2_CUDA(cudaMemsetAsync(dgh, DS2_FILL_BYTE, (size_t)T * B * ND * 3 * H * 2, st));
  return gru_launch<true>(ta, p, geo, st);
}

}  // namespace ds2

using namespace ds2;

extern "C" int ds2_gru_persistent_supported(int B, int H, int ND) { return gru_persist_supported(B, H, ND); }

extern "C" size_t ds2_gru_workspace_bytes(int T, int B, int H, int ND) { return gru_persist_workspace_bytes(T, B, H, ND); }

extern "C" int ds2_gru_persistent_fwd(const float* wx, float scale, const void* u, const float* bw, const float* bu, const int* lens,
                                      void* hs, void* gates, int T, int B, int H, int ND, int reverse_mask, void* workspace,
                                      size_t workspace_bytes, void* stream) {
  DS2_REQUIRE(wx && u && bw && bu && lens && hs && gates, DS2_ERR_BAD_ARG, "ds2_gru_persistent_fwd: null pointer");
  DS2_REQUIRE(T > 0 && B > 0 && H > 0 && (ND == 1 || ND == 2), DS2_ERR_BAD_SHAPE, "ds2_gru_persistent_fwd: bad shape");
  return gru_fwd_persist(wx, scale, u, bw, bu, lens, hs, gates, T, B, H, ND, reverse_mask, workspace, workspace_bytes,
                         static_cast<cudaStream_t>(stream));
}

extern "C" int ds2_gru_persistent_bwd(const float* dout, const void* u, const int* lens, const void* hs, const void* gates, void* dgi,
                                      void* dgh, float* dbw, float* dbu, int T, int B, int H, int ND, int reverse_mask,
                                      void* workspace, size_t workspace_bytes, void* stream) {
  DS2_REQUIRE(dout && u && lens && hs && gates && dgi && dgh, DS2_ERR_BAD_ARG, "ds2_gru_persistent_bwd: null pointer");
  DS2_REQUIRE(T > 0 && B > 0 && H > 0 && (ND == 1 || ND == 2), DS2_ERR_BAD_SHAPE, "ds2_gru_persistent_bwd: bad shape");
  return gru_bwd_persist(dout, u, lens, hs, gates, dgi, dgh, dbw, dbu, T, B, H, ND, reverse_mask, workspace, workspace_bytes,
                         static_cast<cudaStream_t>(stream));
}
