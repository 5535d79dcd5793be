// bf16 x bf16 -> fp32 GEMM on the 5th-gen tensor cores: TMA (128B swizzle) -> smem ring ->
// tcgen05.mma (one elected thread) -> TMEM accumulators (double buffered) -> tcgen05.ld epilogue.
// Persistent: one CTA per SM loops over 128x128 output tiles.
//
//   C[M,N] = alpha * A * B^T (+ bias[n]) (+ beta * C)
//   A: K-major  [M,K] row-major (a_mn=0)  or MN-major [K,M] row-major (a_mn=1)
//   B: K-major  [N,K] row-major (b_mn=0)  or MN-major [K,N] row-major (b_mn=1)
//
// Used for: the batched input projection W.x of every recurrent layer (src/custom_ops.py:60-61
// executed for all timesteps at once), the FC layer (src/deepSpeech_NCHW.py:195-198), conv2 as an
// (explicit-im2col) GEMM, and all batched backward GEMMs (dW, dU, dx).
#include <stdarg.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

namespace ds2 {

static constexpr int BM = 128, BK = 64;
static constexpr int ACC_STAGES = 2;
static constexpr int A_BYTES = BM * BK * 2;
static constexpr int MN_GROUP_BYTES = BK * 128;       // one 64-wide MN group of an MN-major tile: [BK rows][128 B]
// Tile width BN is a template parameter: 256 for wide outputs (a 128x128 tile reads A+B at exactly the
// 128 B/clk shared-memory limit per MMA; 128x256 needs 96 B/clk), 128 / 32 for narrow ones (N = 29, 32).
template <int BN> struct GemmCfg {
  static constexpr int B_BYTES = (BN * BK * 2 < MN_GROUP_BYTES) ? MN_GROUP_BYTES : BN * BK * 2;   // room for one MN-major group
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN == 256) ? 4 : 6;
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
  static constexpr int TMEM_COLS = (ACC_STAGES * BN < 32) ? 32 : ACC_STAGES * BN;
};
static constexpr int GEMM_THREADS = 256;

struct GemmParams {
  void* C;
  long ldc;
  int M, N, K;
  float alpha, beta;
  const float* bias;
  int c_bf16, a_mn, b_mn;
  int num_m_blocks, num_n_blocks;
  int splits, kb_per_split;   // split-K: partial tiles are red.add'ed into a zero-initialised fp32 C
};

template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, GemmParams p) {
  constexpr int NGROUPS = (BN / 64 > 0) ? BN / 64 : 1;   // 64-wide groups of an MN-major B tile
  constexpr int B_BYTES = GemmCfg<BN>::B_BYTES, STAGE_BYTES = GemmCfg<BN>::STAGE_BYTES, STAGES = GemmCfg<BN>::STAGES;
  extern __shared__ uint8_t gemm_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(gemm_smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* full_bar = bars;                         // [STAGES]
  uint64_t* empty_bar = bars + STAGES;               // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;           // [ACC_STAGES]
  uint64_t* tempty_bar = bars + 2 * STAGES + ACC_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 2 * ACC_STAGES);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int mn_tiles = p.num_m_blocks * p.num_n_blocks;
  const int num_tiles = mn_tiles * p.splits;
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(full_bar + i, 1); mbar_init(empty_bar + i, 1); }
    for (int i = 0; i < ACC_STAGES; ++i) { mbar_init(tfull_bar + i, 1); mbar_init(tempty_bar + i, 4); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, GemmCfg<BN>::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one_sync()) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int mn = tile % mn_tiles, split = tile / mn_tiles;
        const int m0 = (mn % p.num_m_blocks) * BM, n0 = (mn / p.num_m_blocks) * BN;
        const int kb0 = split * p.kb_per_split, kb1 = min(num_kb, kb0 + p.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar + stage, phase ^ 1, 100 + stage);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          mbar_arrive_expect_tx(full_bar + stage, A_BYTES + (p.b_mn ? NGROUPS * MN_GROUP_BYTES : BN * BK * 2));
          const int k0 = kb * BK;
          if (!p.a_mn) {
            tma_load_2d(sa, &tmap_a, full_bar + stage, k0, m0);
          } else {
            tma_load_2d(sa, &tmap_a, full_bar + stage, m0, k0);
            tma_load_2d(sa + MN_GROUP_BYTES, &tmap_a, full_bar + stage, m0 + 64, k0);
          }
          if (!p.b_mn) {
            tma_load_2d(sb, &tmap_b, full_bar + stage, k0, n0);
          } else {
#pragma unroll
            for (int g = 0; g < NGROUPS; ++g) tma_load_2d(sb + g * MN_GROUP_BYTES, &tmap_b, full_bar + stage, n0 + 64 * g, k0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = umma_idesc_bf16(BM, BN, p.a_mn, p.b_mn);
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      mbar_wait(tempty_bar + acc, acc_phase ^ 1, 200 + acc);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      const int split = tile / mn_tiles;
      const int kb0 = split * p.kb_per_split, kb1 = min(num_kb, kb0 + p.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar + stage, phase, 300 + stage);
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major: +32 B per K=16 inside the 128B swizzle atom, SBO = 8 rows * 128 B.
            // MN-major: +16 k-rows * 128 B per K=16, LBO = 64-element MN group stride, SBO = 8 k-rows.
            const uint64_t adesc = p.a_mn ? umma_smem_desc_sw128(sa + k * 2048, MN_GROUP_BYTES, 1024)
                                          : umma_smem_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t bdesc = p.b_mn ? umma_smem_desc_sw128(sb + k * 2048, MN_GROUP_BYTES, 1024)
                                          : umma_smem_desc_sw128(sb + k * 32, 16, 1024);
            umma_bf16(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(empty_bar + stage);                 // frees the smem slot when the MMAs retire
          if (kb == kb1 - 1) umma_commit(tfull_bar + acc);   // accumulator ready for the epilogue
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (TMEM -> registers -> global) =====================
    const int wq = warp & 3;      // TMEM lane quarter this warp may access
    int acc = 0; uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int mn = tile % mn_tiles, split = tile / mn_tiles;
      const int m0 = (mn % p.num_m_blocks) * BM, n0 = (mn / p.num_m_blocks) * BN;
      mbar_wait(tfull_bar + acc, acc_phase, 400 + acc);
      tc_fence_after();
      const int row = m0 + wq * 32 + lane;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + acc * BN + c * 32 + (static_cast<uint32_t>(wq * 32) << 16), r);
        tmem_ld_wait();
        const int nbase = n0 + c * 32;
        if (row < p.M && nbase < p.N) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = p.alpha * __uint_as_float(r[j]);
          if (p.bias && split == 0) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (nbase + j < p.N) v[j] += __ldg(p.bias + nbase + j);
          }
          const bool full = nbase + 32 <= p.N;
          if (p.splits > 1) {
            float* crow = static_cast<float*>(p.C) + (long)row * p.ldc + nbase;
            if (full && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)      // one 16-byte reduction instead of four scalar ones
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(crow + j), "f"(v[j]), "f"(v[j + 1]), "f"(v[j + 2]),
                             "f"(v[j + 3]) : "memory");
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) atomicAdd(crow + j, v[j]);
            }
          } else if (p.c_bf16) {
            __nv_bfloat16* crow = static_cast<__nv_bfloat16*>(p.C) + (long)row * p.ldc + nbase;
            if (p.beta != 0.f) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) v[j] += p.beta * __bfloat162float(crow[j]);
            }
            if (full && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                uint4 pk;
                __nv_bfloat162 b0 = __floats2bfloat162_rn(v[j], v[j + 1]), b1 = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
                __nv_bfloat162 b2 = __floats2bfloat162_rn(v[j + 4], v[j + 5]), b3 = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
                pk.x = *reinterpret_cast<uint32_t*>(&b0); pk.y = *reinterpret_cast<uint32_t*>(&b1);
                pk.z = *reinterpret_cast<uint32_t*>(&b2); pk.w = *reinterpret_cast<uint32_t*>(&b3);
                *reinterpret_cast<uint4*>(crow + j) = pk;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) crow[j] = __float2bfloat16(v[j]);
            }
          } else {
            float* crow = static_cast<float*>(p.C) + (long)row * p.ldc + nbase;
            if (p.beta != 0.f) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) v[j] += p.beta * crow[j];
            }
            if (full && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(crow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) crow[j] = v[j];
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar + acc);
      if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, GemmCfg<BN>::TMEM_COLS);
  }
}

// ------------------------------------------------------------------ CTA-pair variant (tcgen05 cta_group::2)
// A 128x256 single-CTA tile writes (TMA) and reads (MMA) ~180 B/clk of shared memory against a 128 B/clk port: 70 % of
// peak.  Two CTAs of a cluster share one 256x256 tile instead: each loads its 128 rows of A and 128 of the 256 rows of
// B (32 KB per stage instead of 48), the leader issues one M=256 MMA that reads both halves of B, and each CTA's TMEM
// receives its own 128 rows of D.  All TMA transactions of a stage complete on the LEADER's full barrier; slot release
// and accumulator hand-over are multicast tcgen05.commit's to both CTAs; the peer's epilogue warps arrive remotely on
// the leader's accumulator-free barrier.
static constexpr int G2_STAGES = 6;
static constexpr int G2_HALF_BYTES = 128 * BK * 2;                    // 16 KB: one operand half of one CTA
static constexpr int G2_STAGE_BYTES = 2 * G2_HALF_BYTES;
static constexpr int G2_SMEM = G2_STAGES * G2_STAGE_BYTES + 1024 + 256;
static constexpr int G2_BN = 256;

__device__ __forceinline__ void tma_load_2d_pair(uint32_t smem_dst, const void* tmap, uint32_t leader_bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(leader_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {      // arrives on `bar` of BOTH CTAs
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((unsigned short)3) : "memory");
}

__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, GemmParams p) {
  extern __shared__ uint8_t gemm2_smem_raw[];
  const uint32_t sbase = (smem_u32(gemm2_smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = sbase + G2_STAGES * G2_STAGE_BYTES;
  const uint32_t full_bar = bars, empty_bar = bars + 8 * G2_STAGES, tfull_bar = bars + 16 * G2_STAGES, tempty_bar = tfull_bar + 16;
  const uint32_t slot_addr = tempty_bar + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gemm2_smem_raw + (slot_addr - smem_u32(gemm2_smem_raw)));
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int mn_tiles = p.num_m_blocks * p.num_n_blocks;
  const int num_tiles = mn_tiles * p.splits;
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < G2_STAGES; ++i) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full_bar + 8 * i), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(empty_bar + 8 * i), "r"(1));
    }
    for (int i = 0; i < ACC_STAGES; ++i) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tfull_bar + 8 * i), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tempty_bar + 8 * i), "r"(8));   // 4 epilogue warps x 2 CTAs
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_addr), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs; every transaction lands on the leader's full barrier) =====================
    if (elect_one_sync()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int mn = tile % mn_tiles, split = tile / mn_tiles;
        const int m0 = (mn % p.num_m_blocks) * 256 + 128 * (int)rank, n0 = (mn / p.num_m_blocks) * G2_BN + 128 * (int)rank;
        const int kb0 = split * p.kb_per_split, kb1 = min(num_kb, kb0 + p.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar + 8 * stage, phase ^ 1, 100 + stage);
          const uint32_t sa = sbase + stage * G2_STAGE_BYTES, sb = sa + G2_HALF_BYTES;
          const uint32_t lbar = (full_bar + 8 * stage) & 0xFEFFFFFFu;       // same barrier in the cluster's CTA 0
          if (leader) mbar_arrive_expect_tx(full_bar + 8 * stage, 2 * G2_STAGE_BYTES);
          const int k0 = kb * BK;
          if (!p.a_mn) {
            tma_load_2d_pair(sa, &tmap_a, lbar, k0, m0);
          } else {
            tma_load_2d_pair(sa, &tmap_a, lbar, m0, k0);
            tma_load_2d_pair(sa + MN_GROUP_BYTES, &tmap_a, lbar, m0 + 64, k0);
          }
          if (!p.b_mn) {
            tma_load_2d_pair(sb, &tmap_b, lbar, k0, n0);
          } else {
            tma_load_2d_pair(sb, &tmap_b, lbar, n0, k0);
            tma_load_2d_pair(sb + MN_GROUP_BYTES, &tmap_b, lbar, n0 + 64, k0);
          }
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1 && leader) {
    // ===================== MMA issuer (leader CTA only) =====================
    const uint32_t idesc = umma_idesc_bf16(256, G2_BN, p.a_mn, p.b_mn);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1, 200 + acc);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * G2_BN;
      const int split = tile / mn_tiles;
      const int kb0 = split * p.kb_per_split, kb1 = min(num_kb, kb0 + p.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar + 8 * stage, phase, 300 + stage);
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t sa = sbase + stage * G2_STAGE_BYTES, sb = sa + G2_HALF_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t adesc = p.a_mn ? umma_smem_desc_sw128(sa + k * 2048, MN_GROUP_BYTES, 1024)
                                          : umma_smem_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t bdesc = p.b_mn ? umma_smem_desc_sw128(sb + k * 2048, MN_GROUP_BYTES, 1024)
                                          : umma_smem_desc_sw128(sb + k * 32, 16, 1024);
            umma_bf16_pair(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit_pair(empty_bar + 8 * stage);
          if (kb == kb1 - 1) umma_commit_pair(tfull_bar + 8 * acc);
        }
        __syncwarp();
        if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: this CTA's 128 rows of the tile =====================
    const int wq = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t leader_tempty;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(leader_tempty) : "r"(tempty_bar), "r"(0));
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int mn = tile % mn_tiles, split = tile / mn_tiles;
      const int m0 = (mn % p.num_m_blocks) * 256 + 128 * (int)rank, n0 = (mn / p.num_m_blocks) * G2_BN;
      mbar_wait(tfull_bar + 8 * acc, acc_phase, 400 + acc);
      tc_fence_after();
      const int row = m0 + wq * 32 + lane;
#pragma unroll 1
      for (int c = 0; c < G2_BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + acc * G2_BN + c * 32 + (static_cast<uint32_t>(wq * 32) << 16), r);
        tmem_ld_wait();
        const int nbase = n0 + c * 32;
        if (row < p.M && nbase < p.N) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = p.alpha * __uint_as_float(r[j]);
          if (p.bias && split == 0) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (nbase + j < p.N) v[j] += __ldg(p.bias + nbase + j);
          }
          const bool full = nbase + 32 <= p.N;
          if (p.splits > 1) {
            float* crow = static_cast<float*>(p.C) + (long)row * p.ldc + nbase;
            if (full && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)      // one 16-byte reduction instead of four scalar ones
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(crow + j), "f"(v[j]), "f"(v[j + 1]), "f"(v[j + 2]),
                             "f"(v[j + 3]) : "memory");
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) atomicAdd(crow + j, v[j]);
            }
          } else if (p.c_bf16) {
            __nv_bfloat16* crow = static_cast<__nv_bfloat16*>(p.C) + (long)row * p.ldc + nbase;
            if (p.beta != 0.f) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) v[j] += p.beta * __bfloat162float(crow[j]);
            }
            if (full && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                uint4 pk;
                __nv_bfloat162 b0 = __floats2bfloat162_rn(v[j], v[j + 1]), b1 = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
                __nv_bfloat162 b2 = __floats2bfloat162_rn(v[j + 4], v[j + 5]), b3 = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
                pk.x = *reinterpret_cast<uint32_t*>(&b0); pk.y = *reinterpret_cast<uint32_t*>(&b1);
                pk.z = *reinterpret_cast<uint32_t*>(&b2); pk.w = *reinterpret_cast<uint32_t*>(&b3);
                *reinterpret_cast<uint4*>(crow + j) = pk;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) crow[j] = __float2bfloat16(v[j]);
            }
          } else {
            float* crow = static_cast<float*>(p.C) + (long)row * p.ldc + nbase;
            if (p.beta != 0.f) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) v[j] += p.beta * crow[j];
            }
            if (full && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(crow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (nbase + j < p.N) crow[j] = v[j];
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(leader_tempty + 8 * acc) : "memory");
      if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t stride_outer_elems,
                      uint32_t box_inner, uint32_t box_outer) {
  PFN_encodeTiled enc = get_encode_fn();
  DS2_REQUIRE(enc != nullptr, DS2_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available from the driver");
  DS2_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, DS2_ERR_BAD_ARG, "TMA operand base must be 16-byte aligned");
  DS2_REQUIRE((stride_outer_elems * 2) % 16 == 0, DS2_ERR_BAD_SHAPE, "TMA operand row pitch (%llu elements) must be a multiple of 8",
              (unsigned long long)stride_outer_elems);
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {stride_outer_elems * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DS2_REQUIRE(r == CUDA_SUCCESS, DS2_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu outer=%llu pitch=%llu)",
              (int)r, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)stride_outer_elems);
  return DS2_OK;
}

int gemm_bf16_launch(const void* A, long lda, int a_mn, const void* B, long ldb, int b_mn, void* C, int c_dtype, long ldc,
                     int M, int N, int K, float alpha, float beta, const float* bias, int splits, cudaStream_t st) {
  DS2_REQUIRE(A && B && C, DS2_ERR_BAD_ARG, "ds2_gemm_bf16: null pointer");
  DS2_REQUIRE(M > 0 && N > 0 && K > 0, DS2_ERR_BAD_SHAPE, "ds2_gemm_bf16: bad shape M=%d N=%d K=%d", M, N, K);
  DS2_REQUIRE(c_dtype == DS2_F32 || c_dtype == DS2_BF16, DS2_ERR_BAD_ARG, "ds2_gemm_bf16: bad c_dtype %d", c_dtype);
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn) rc = make_tmap_2d_bf16(&ta, A, K, M, lda, BK, BM);
  else rc = make_tmap_2d_bf16(&ta, A, M, K, lda, 64, BK);
  if (rc) return rc;
  const int bn = (N <= 32) ? 32 : (N <= 128 ? 128 : 256);
  // wide outputs with at least two row blocks: CTA pairs on 256x256 tiles (DS2_GEMM_NO_PAIR=1: single-CTA 128x256)
  const bool no_pair = env().gemm_no_pair != 0;
  const bool pair = bn == 256 && M > 128 && !no_pair && num_sms() >= 2;
  if (!b_mn) rc = make_tmap_2d_bf16(&tb, B, K, N, ldb, BK, pair ? 128 : bn);
  else rc = make_tmap_2d_bf16(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  static unsigned long long attr_done_devs = 0;
  if (first_use_on_device(&attr_done_devs)) {
    DS2_CUDA(cudaFuncSetAttribute(gemm_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G2_SMEM));
    DS2_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<32>::SMEM));
    DS2_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<128>::SMEM));
    DS2_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<256>::SMEM));
  }
  GemmParams p;
  p.C = C; p.ldc = ldc; p.M = M; p.N = N; p.K = K; p.alpha = alpha; p.beta = beta; p.bias = bias;
  p.c_bf16 = (c_dtype == DS2_BF16); p.a_mn = a_mn; p.b_mn = b_mn;
  p.num_m_blocks = pair ? (M + 255) / 256 : (M + BM - 1) / BM;
  p.num_n_blocks = (N + bn - 1) / bn;
  const int num_kb = (K + BK - 1) / BK;
  if (splits < 1) splits = 1;
  if (splits > num_kb) splits = num_kb;
  p.kb_per_split = (num_kb + splits - 1) / splits;
  p.splits = (num_kb + p.kb_per_split - 1) / p.kb_per_split;     // no empty split
  if (p.splits > 1) {
    DS2_REQUIRE(c_dtype == DS2_F32 && beta == 0.f, DS2_ERR_BAD_ARG, "ds2_gemm_bf16: split-K needs an fp32 C and beta = 0");
    DS2_CUDA(cudaMemset2DAsync(C, ldc * sizeof(float), 0, (size_t)N * sizeof(float), M, st));
  }
  const int tiles = p.num_m_blocks * p.num_n_blocks * p.splits;
  if (pair) {
    const int clusters = tiles < num_sms() / 2 ? tiles : num_sms() / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = G2_SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 2;
    attr.val.clusterDim.y = 1;
    attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_pair_kernel, ta, tb, p);
    if (e != cudaSuccess) return cuda_fail(e, "gemm_pair_kernel");
    count_launch();
    return DS2_OK;
  }
  const int grid = tiles < num_sms() ? tiles : num_sms();
  if (bn == 32) gemm_tcgen05_kernel<32><<<grid, GEMM_THREADS, GemmCfg<32>::SMEM, st>>>(ta, tb, p);
  else if (bn == 128) gemm_tcgen05_kernel<128><<<grid, GEMM_THREADS, GemmCfg<128>::SMEM, st>>>(ta, tb, p);
  else gemm_tcgen05_kernel<256><<<grid, GEMM_THREADS, GemmCfg<256>::SMEM, st>>>(ta, tb, p);
  DS2_LAUNCH_CHECK("gemm_tcgen05_kernel");
  count_launch();
  return DS2_OK;
}

}  // namespace ds2

extern "C" int ds2_gemm_bf16(const void* A, long lda, int a_mn_major, const void* B, long ldb, int b_mn_major, void* C,
                             int c_dtype, long ldc, int M, int N, int K, float alpha, float beta, const float* bias,
                             void* stream) {
  return ds2::gemm_bf16_launch(A, lda, a_mn_major, B, ldb, b_mn_major, C, c_dtype, ldc, M, N, K, alpha, beta, bias, 1,
                               static_cast<cudaStream_t>(stream));
}

extern "C" int ds2_gemm_bf16_splitk(const void* A, long lda, int a_mn_major, const void* B, long ldb, int b_mn_major, float* C,
                                    long ldc, int M, int N, int K, float alpha, const float* bias, int splits, void* stream) {
  if (splits <= 0) {
    // auto: the split count that minimises tile waves per unit of K, ceil(tiles*s / units) / s, a "unit" being what runs
    // one tile at a time (an SM, or a CTA pair for the 256x256 tiles).  dU [1760x1760] has 49 pair tiles for 74 pairs
    // (one wave at 66 %); three K splits make it two full waves of a third of K each.  A split keeps >= 8 K-blocks and
    // pays for the red.add epilogue and the zero fill: taken only for a gain of DS2_GEMM_SPLIT_GAIN percent (default 10).
    const int bn = (N <= 32) ? 32 : (N <= 128 ? 128 : 256);
    const bool pair = bn == 256 && M > 128 && !ds2::env().gemm_no_pair;
    const int tiles = pair ? ((M + 255) / 256) * ((N + 255) / 256) : ((M + ds2::BM - 1) / ds2::BM) * ((N + bn - 1) / bn);
    const int units = pair ? ds2::num_sms() / 2 : ds2::num_sms();
    const int num_kb = (K + ds2::BK - 1) / ds2::BK;
    static const double gain = getenv("DS2_GEMM_SPLIT_GAIN") ? atof(getenv("DS2_GEMM_SPLIT_GAIN")) / 100.0 : 0.10;
    splits = 1;
    double best = (double)((tiles + units - 1) / units);
    for (int sp = 2; sp <= 32 && num_kb / sp >= 8; ++sp) {
      const double cost = (double)(((long)tiles * sp + units - 1) / units) / sp;
      if (cost < best * (1.0 - gain)) {
        best = cost;
        splits = sp;
      }
    }
  }
  return ds2::gemm_bf16_launch(A, lda, a_mn_major, B, ldb, b_mn_major, C, DS2_F32, ldc, M, N, K, alpha, 0.f, bias, splits,
                               static_cast<cudaStream_t>(stream));
}
