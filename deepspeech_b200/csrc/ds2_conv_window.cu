// Implicit-GEMM 2-D convolution for 32-channel layers with unit stride along frequency (conv2 of the reference:
// tf.nn.conv2d at src/deepSpeech_NCHW.py:142-151 / src/deepSpeech.py:143-149, filter [10,5,32,32], strides (2,1)),
// forward, data gradient and weight gradient, WITHOUT a patch matrix (the explicit im2col route writes and re-reads
// 2.8 GB per pass at the reference's batch).
//
// The trick is the tensor map.  With channel-last bf16 activations a[R rows = (b,t)][P positions f][32 ch], the
// K-major A row of output position f for filter row kt is the contiguous run a[row][f .. f+KF-1][0..31]: windows of
// neighbouring positions OVERLAP, shifted by one position = 64 bytes.  A 3-D TMA view {64 elements, P, R} with strides
// {64 B, row pitch} describes exactly that: box {64, 128, 1} at (0, f0 + 2c, row) lands 128 windows of 64 elements
// (two filter taps x 32 channels) as a 128-byte-swizzled K-major tile, one per output position.  A filter row of KF
// taps is KF/2 such chunks plus a half chunk (32 elements: only two of the four K=16 MMA steps are issued).
// Positions beyond the row are out of bounds for the map and arrive as zeros.
// A row of conv2 has 75 output positions, 79 for the data gradient: one row per 128-row MMA tile would idle 40 % of the
// tensor pipe.  So a tile PACKS rows: the box is {64, PB positions, RB rows} with PB * RB <= 128 (75 -> 3 passes of 25
// positions x 5 rows = 125 MMA rows; 79 -> 5 passes of 16 x 8 = 128), landing as PB*RB consecutive 128-byte rows.
//   forward : tile = (b, group of RB output rows, position pass), M = 128, N = 32 filters, K-blocks = KT x chunks.
//             Output rows to, to+1, .. read source rows st apart: one tensor map per residue of the source row modulo st
//             (base shifted by that many rows, row pitch st times the tensor's), so that a box takes RB consecutive rows
//             of its map.  Weights Wt[co][(kt,kf,ci)].
//   dgrad   : tile = (b, residue of the input row t1 modulo st, group of RB such rows, pass); source = dy zero-padded by
//             KF-1 positions on both sides AND by zero rows above and below every batch item (the rows to = (t1-kt)/st
//             outside [0,To) then read zeros, whichever item's memory they fall into); K-blocks over the kt of the same
//             residue, source rows consecutive.  Weights Wd[ci][(kt,KF-1-kf,co)].
//   wgrad   : dW[(kt,kf,ci)][co] = sum over tiles of window^T . dy : the same boxes read as MN-major A operands
//             (M = 64 window elements per box, two boxes per MMA block), the packed dy rows as the MN-major B operand
//             (contraction over the PB*RB packed rows; the stage buffers are zeroed once, rows the boxes never write
//             stay zero), all KT*KF*32 x 32 partial sums of a CTA resident in TMEM (15 blocks x 32 columns) until it has
//             run out of tiles, then red.add into dW.
#include "common.cuh"

namespace ds2 {

static constexpr int CW_THREADS = 256;
static constexpr int CW_C = 32;                  // channels in and out
static constexpr int CW_BOX = 128 * 128;         // one window box: 128 positions x 64 elements bf16
static constexpr int CW_WBOX = 32 * 128;         // weight tile: 32 rows x 64 elements
static constexpr int CW_STAGES = 6;              // window boxes in flight; the filters stay resident behind them
static constexpr int CW_STAGE_BYTES = CW_BOX;
static constexpr int CW_MAX_WBOXES = 32;         // KT x chunks per filter row
static constexpr int CW_W_OFF = CW_STAGES * CW_STAGE_BYTES;
static constexpr int CW_BUF_BYTES = CW_W_OFF + CW_MAX_WBOXES * CW_WBOX;
static constexpr int CW_SMEM = CW_BUF_BYTES + 1024 + 256 + 1024;   // + the epilogue warps' partial channel sums
static constexpr int CWG_STAGES = 5;             // wgrad: two window boxes per stage; the tile's dy box in one of two slots
static constexpr int CWG_STAGE_BYTES = 2 * CW_BOX;
static constexpr int CWG_DY_OFF = CWG_STAGES * CWG_STAGE_BYTES;
static constexpr int CWG_BUF_BYTES = CWG_DY_OFF + 2 * CW_BOX;
static constexpr int CWG_SMEM = CWG_BUF_BYTES + 1024 + 256;

static constexpr int CW_MAXST = 4;               // row packing needs one source map per residue modulo the time stride
struct WinMaps {
  CUtensorMap m[CW_MAXST];
};
struct WinParams {
  float* out;            // fwd: y [B,Rout,Pout,32] ; dgrad: da [B,Rout,Pout,32] ; wgrad: dW [KT*KF*32][32]
  const float* bias;     // fwd only
  double* sums;          // fwd, optional: [64] per-channel sum and sum of squares of the output (batch-norm statistics), zeroed by the host
  int n_tiles;           // B * classes * G * np
  int Rout, Rsrc;        // rows per batch item of the output / of the source (dgrad, wgrad dy: padded rows per item)
  int Pout;              // valid positions per output row
  int PB, RB, np, G;     // packing: positions and rows per box, position passes, row groups per batch item (and class)
  int ncls;              // dgrad: st residue classes of the output row; otherwise 1
  int nmaps;             // fwd / wgrad source maps: st (one per residue) or 1 (RB == 1: plain rows)
  int ptop;              // dgrad / wgrad: zero rows above every batch item in the padded dy
  int Rdy;               // wgrad: rows per batch item of the padded dy
  int KT, st;
  int nfull, half;       // 64-element chunks per filter row, and whether a 32-element one follows
  int kstride;           // K elements per filter row (KF * 32)
  int win0;              // first window position of output position 0 (wgrad: position of dy's first valid column)
  int dgrad;
};

// tile index -> batch item, residue class, row group, position pass (pass fastest: the passes of a row group read the
// same source rows back to back)
struct TileId {
  int b, cls, g, q;
};
__device__ __forceinline__ TileId tile_decode(const WinParams& p, int tile) {
  TileId t;
  t.q = tile % p.np;
  tile /= p.np;
  t.g = tile % p.G;
  tile /= p.G;
  t.cls = tile % p.ncls;
  t.b = tile / p.ncls;
  return t;
}
// source map and first row coordinate of filter row kt for the tile's first packed row; false: kt not in the tile's class
__device__ __forceinline__ bool kt_source(const WinParams& p, const TileId& t, int kt, int* map, int* row) {
  if (!p.dgrad) {
    const int r0 = t.b * p.Rsrc + t.g * p.RB * p.st + kt;
    if (p.nmaps == 1) { *map = 0; *row = r0; }
    else { *map = r0 % p.st; *row = r0 / p.st; }
    return true;
  }
  if (kt % p.st != t.cls) return false;
  *map = 0;
  *row = t.b * p.Rsrc + p.ptop + (t.cls - kt) / p.st + t.g * p.RB;     // (cls - kt) is a non-positive multiple of st
  return true;
}

__global__ void __launch_bounds__(CW_THREADS, 1)
conv_window_kernel(const __grid_constant__ WinMaps maps, const __grid_constant__ CUtensorMap tmap_w, WinParams p) {
  extern __shared__ __align__(1024) uint8_t cw_smem_raw[];
  const uint32_t sbase = (smem_u32(cw_smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = sbase + CW_BUF_BYTES;
  const uint32_t full_bar = bars, empty_bar = bars + 8 * CW_STAGES, tfull_bar = bars + 16 * CW_STAGES,
                 tempty_bar = tfull_bar + 16, w_bar = tempty_bar + 16;
  const uint32_t slot_addr = w_bar + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(cw_smem_raw + (slot_addr - smem_u32(cw_smem_raw)));
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int nch = p.nfull + p.half;

  const uint32_t box_bytes = (uint32_t)p.PB * p.RB * 128u;
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.nmaps; ++i) tma_prefetch_desc(&maps.m[i]);
    tma_prefetch_desc(&tmap_w);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < CW_STAGES; ++i) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full_bar + 8 * i), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(empty_bar + 8 * i), "r"(1));
    }
    for (int i = 0; i < 2; ++i) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tfull_bar + 8 * i), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tempty_bar + 8 * i), "r"(4));
    }
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(w_bar), "r"(1));
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one_sync()) {
      // all filter boxes once: [32 filters x 64 window elements] per (filter row, chunk)
      mbar_arrive_expect_tx(w_bar, (uint32_t)(p.KT * nch) * CW_WBOX);
      for (int i = 0; i < p.KT * nch; ++i) {
        const int kt = i / nch, c = i - kt * nch;
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
            ::"r"(sbase + CW_W_OFF + i * CW_WBOX), "l"(reinterpret_cast<uint64_t>(&tmap_w)), "r"(w_bar),
            "r"(kt * p.kstride + c * 64), "r"(0)
            : "memory");
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        const TileId t = tile_decode(p, tile);
        for (int kt = 0; kt < p.KT; ++kt) {
          int mi, src_row;
          if (!kt_source(p, t, kt, &mi, &src_row)) continue;
          for (int c = 0; c < nch; ++c) {
            mbar_wait(empty_bar + 8 * stage, phase ^ 1, 100 + stage);
            const uint32_t sa = sbase + stage * CW_STAGE_BYTES;
            mbar_arrive_expect_tx(full_bar + 8 * stage, box_bytes);
            asm volatile(
                "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                ::"r"(sa), "l"(reinterpret_cast<uint64_t>(&maps.m[mi])), "r"(full_bar + 8 * stage), "r"(0),
                "r"(p.win0 + t.q * p.PB + 2 * c), "r"(src_row)
                : "memory");
            if (++stage == CW_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = umma_idesc_bf16(128, CW_C, 0, 0);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    mbar_wait(w_bar, 0, 250);          // the filters are resident
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const TileId t = tile_decode(p, tile);
      mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1, 200 + acc);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * CW_C;
      bool first = true;
      for (int kt = 0; kt < p.KT; ++kt) {
        int mi, src_row;
        if (!kt_source(p, t, kt, &mi, &src_row)) continue;
        for (int c = 0; c < nch; ++c) {
          mbar_wait(full_bar + 8 * stage, phase, 300 + stage);
          tc_fence_after();
          if (elect_one_sync()) {
            const uint32_t sa = sbase + stage * CW_STAGE_BYTES;
            const uint32_t sb = sbase + CW_W_OFF + (kt * nch + c) * CW_WBOX;
            const int ksteps = c < p.nfull ? 4 : 2;
            for (int k = 0; k < ksteps; ++k) {
              umma_bf16(d_tmem, umma_smem_desc_sw128(sa + k * 32, 16, 1024), umma_smem_desc_sw128(sb + k * 32, 16, 1024), idesc,
                        (first && k == 0) ? 0u : 1u);
            }
            umma_commit(empty_bar + 8 * stage);
          }
          __syncwarp();
          first = false;
          if (++stage == CW_STAGES) { stage = 0; phase ^= 1; }
        }
      }
      if (elect_one_sync()) umma_commit(tfull_bar + 8 * acc);   // also fires when the tile had no K-block at all
      __syncwarp();
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> 128 contiguous bytes per output position =====================
    const int wq = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    float bias[CW_C];
#pragma unroll
    for (int j = 0; j < CW_C; ++j) bias[j] = p.bias ? __ldg(p.bias + j) : 0.f;
    // accumulator row of this thread -> packed row and position inside the box
    const int mrow = wq * 32 + lane;
    const int pi = mrow / p.PB, pp = mrow - pi * p.PB;
    float s1[CW_C], s2[CW_C];            // running per-channel sums of this thread's outputs (p.sums)
#pragma unroll
    for (int j = 0; j < CW_C; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const TileId t = tile_decode(p, tile);
      bool any = false;
      for (int kt = 0; kt < p.KT; ++kt) {
        int mi, src_row;
        any |= kt_source(p, t, kt, &mi, &src_row);
      }
      mbar_wait(tfull_bar + 8 * acc, acc_phase, 400 + acc);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + acc * CW_C + (static_cast<uint32_t>(wq * 32) << 16), v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar + 8 * acc);
      const int pos = t.q * p.PB + pp;
      const int r = p.dgrad ? t.cls + p.st * (t.g * p.RB + pi) : t.g * p.RB + pi;
      if (pi < p.RB && pos < p.Pout && r < p.Rout) {
        float4* dst = reinterpret_cast<float4*>(p.out + (((size_t)t.b * p.Rout + r) * p.Pout + pos) * CW_C);
#pragma unroll
        for (int j = 0; j < CW_C; j += 4) {
          float4 o;
          o.x = (any ? __uint_as_float(v[j]) : 0.f) + bias[j];
          o.y = (any ? __uint_as_float(v[j + 1]) : 0.f) + bias[j + 1];
          o.z = (any ? __uint_as_float(v[j + 2]) : 0.f) + bias[j + 2];
          o.w = (any ? __uint_as_float(v[j + 3]) : 0.f) + bias[j + 3];
          dst[j / 4] = o;
          if (p.sums) {
            s1[j] += o.x; s1[j + 1] += o.y; s1[j + 2] += o.z; s1[j + 3] += o.w;
            s2[j] = fmaf(o.x, o.x, s2[j]); s2[j + 1] = fmaf(o.y, o.y, s2[j + 1]);
            s2[j + 2] = fmaf(o.z, o.z, s2[j + 2]); s2[j + 3] = fmaf(o.w, o.w, s2[j + 3]);
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (p.sums) {
      float* red = reinterpret_cast<float*>(cw_smem_raw + (slot_addr + 16 - smem_u32(cw_smem_raw)));
#pragma unroll
      for (int j = 0; j < CW_C; ++j) {
        s1[j] = warp_sum(s1[j]);
        s2[j] = warp_sum(s2[j]);
      }
      if (lane == 0) {
#pragma unroll
        for (int j = 0; j < CW_C; ++j) {
          red[wq * 64 + j] = s1[j];
          red[wq * 64 + 32 + j] = s2[j];
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");      // the four epilogue warps
      const int e = threadIdx.x - 128;
      if (e >= 0 && e < 64) atomicAdd(p.sums + e, (double)red[e] + (double)red[64 + e] + (double)red[128 + e] + (double)red[192 + e]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

// ------------------------------------------------------------------ weight gradient
__global__ void __launch_bounds__(CW_THREADS, 1)
conv_window_wgrad_kernel(const __grid_constant__ WinMaps maps, const __grid_constant__ CUtensorMap tmap_dy, WinParams p) {
  extern __shared__ __align__(1024) uint8_t cwg_smem_raw[];
  const uint32_t sbase = (smem_u32(cwg_smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = sbase + CWG_BUF_BYTES;
  const uint32_t full_bar = bars, empty_bar = bars + 8 * CWG_STAGES, done_bar = bars + 16 * CWG_STAGES;
  const uint32_t dyfull_bar = done_bar + 8, dyempty_bar = done_bar + 24;
  const uint32_t slot_addr = done_bar + 40;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(cwg_smem_raw + (slot_addr - smem_u32(cwg_smem_raw)));
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int nch = p.nfull + p.half;
  const int ngroups = p.KT * nch;                  // 64-wide groups of window elements; two per MMA block
  const int nblocks = (ngroups + 1) / 2;

  const uint32_t box_bytes = (uint32_t)p.PB * p.RB * 128u;
  const int ksteps = (p.PB * p.RB + 15) / 16;       // contraction over the packed rows, 16 per MMA
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.nmaps; ++i) tma_prefetch_desc(&maps.m[i]);
    tma_prefetch_desc(&tmap_dy);
  }
  // rows of a stage buffer that no box writes are part of the contraction: zero once
  for (uint32_t o = threadIdx.x * 16u; o < (uint32_t)CWG_BUF_BYTES; o += CW_THREADS * 16u)
    sts_v4(sbase + o, make_uint4(0u, 0u, 0u, 0u));
  fence_proxy_async();
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < CWG_STAGES; ++i) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full_bar + 8 * i), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(empty_bar + 8 * i), "r"(1));
    }
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(done_bar), "r"(1));
    for (int i = 0; i < 2; ++i) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dyfull_bar + 8 * i), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dyempty_bar + 8 * i), "r"(1));
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const bool has_tiles = (int)blockIdx.x < p.n_tiles;

  if (warp == 0) {
    if (elect_one_sync()) {
      int stage = 0;
      uint32_t phase = 0;
      int dslot = 0;
      uint32_t dphase = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        const TileId t = tile_decode(p, tile);
        // the tile's dy rows: loaded once, the B operand of all its blocks
        mbar_wait(dyempty_bar + 8 * dslot, dphase ^ 1, 150 + dslot);
        mbar_arrive_expect_tx(dyfull_bar + 8 * dslot, box_bytes);
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
            ::"r"(sbase + CWG_DY_OFF + dslot * CW_BOX), "l"(reinterpret_cast<uint64_t>(&tmap_dy)), "r"(dyfull_bar + 8 * dslot), "r"(0),
            "r"(p.win0 + t.q * p.PB), "r"(t.b * p.Rdy + p.ptop + t.g * p.RB)
            : "memory");
        if (++dslot == 2) { dslot = 0; dphase ^= 1; }
        for (int blk = 0; blk < nblocks; ++blk) {
          mbar_wait(empty_bar + 8 * stage, phase ^ 1, 100 + stage);
          const uint32_t sa = sbase + stage * CWG_STAGE_BYTES;
          const int ng = min(2, ngroups - 2 * blk);
          mbar_arrive_expect_tx(full_bar + 8 * stage, ng * box_bytes);
          for (int g = 0; g < ng; ++g) {
            const int gi = 2 * blk + g, kt = gi / nch, c = gi - kt * nch;
            int mi, src_row;
            kt_source(p, t, kt, &mi, &src_row);
            asm volatile(
                "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                ::"r"(sa + g * CW_BOX), "l"(reinterpret_cast<uint64_t>(&maps.m[mi])), "r"(full_bar + 8 * stage), "r"(0),
                "r"(t.q * p.PB + 2 * c), "r"(src_row)
                : "memory");
          }
          if (++stage == CWG_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc_bf16(128, CW_C, 1, 1);
    int stage = 0;
    uint32_t phase = 0;
    bool first_tile = true;
    int dslot = 0;
    uint32_t dphase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      mbar_wait(dyfull_bar + 8 * dslot, dphase, 350 + dslot);
      const uint32_t sb = sbase + CWG_DY_OFF + dslot * CW_BOX;
      for (int blk = 0; blk < nblocks; ++blk) {
        mbar_wait(full_bar + 8 * stage, phase, 300 + stage);
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t sa = sbase + stage * CWG_STAGE_BYTES;
          for (int k = 0; k < ksteps; ++k) {
            // contraction over the packed (row, position) pairs of the tile: 16 of them (16 x 128 B) per K step
            umma_bf16(tmem_base + blk * CW_C, umma_smem_desc_sw128(sa + k * 2048, CW_BOX, 1024),
                      umma_smem_desc_sw128(sb + k * 2048, CW_BOX, 1024), idesc, (first_tile && k == 0) ? 0u : 1u);
          }
          umma_commit(empty_bar + 8 * stage);
        }
        __syncwarp();
        if (++stage == CWG_STAGES) { stage = 0; phase ^= 1; }
      }
      first_tile = false;
      if (elect_one_sync()) umma_commit(dyempty_bar + 8 * dslot);     // the tile's MMAs have read its dy slot
      __syncwarp();
      if (++dslot == 2) { dslot = 0; dphase ^= 1; }
    }
    if (elect_one_sync()) umma_commit(done_bar);
    __syncwarp();
  } else if (warp >= 4 && has_tiles) {
    // ===================== after the last tile: TMEM partial sums -> red.add into dW =====================
    const int wq = warp & 3;
    mbar_wait(done_bar, 0, 400);
    tc_fence_after();
    const int ml = wq * 32 + lane;                 // row of the MMA block = window element
    const int g = ml >> 6, e = ml & 63;
    for (int blk = 0; blk < nblocks; ++blk) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + blk * CW_C + (static_cast<uint32_t>(wq * 32) << 16), v);
      tmem_ld_wait();
      const int gi = 2 * blk + g;
      if (gi < ngroups) {
        const int kt = gi / nch, c = gi - kt * nch;
        if (c < p.nfull || e < 32) {
          float* dst = p.out + ((size_t)kt * p.kstride + c * 64 + e) * CW_C;
#pragma unroll
          for (int j = 0; j < CW_C; ++j) atomicAdd(dst + j, __uint_as_float(v[j]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------ small helpers
// dy fp32 [B, To, P, 32] -> bf16 [B, Rdy, P + 2*pad, 32]: zero borders of `pad` positions, `ptop` zero rows above the To
// rows of every batch item and zero rows below them up to Rdy (the data gradient's and weight gradient's source)
__global__ void conv_pad_cast_kernel(const float* __restrict__ dy, __nv_bfloat16* __restrict__ out, int B, int To, int P, int pad,
                                     int ptop, int Rdy) {
  const int Pp = P + 2 * pad;
  const long n4 = (long)B * Rdy * Pp * (CW_C / 4);
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % (CW_C / 4));
    const long rp = i / (CW_C / 4);
    const int pp = (int)(rp % Pp);
    const long rr = rp / Pp;
    const int b = (int)(rr / Rdy), r = (int)(rr - (long)b * Rdy) - ptop;
    const int pos = pp - pad;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (pos >= 0 && pos < P && r >= 0 && r < To) v = __ldg(reinterpret_cast<const float4*>(dy + ((((long)b * To + r) * P + pos) * CW_C)) + c4);
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b2 = __floats2bfloat162_rn(v.z, v.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&a);
    pk.y = *reinterpret_cast<uint32_t*>(&b2);
    *reinterpret_cast<uint2*>(out + (rp * CW_C) + c4 * 4) = pk;
  }
}

// HWIO fp32 w[kt][kf][ci][co] -> wt[co][(kt,kf,ci)] (forward) and wd[ci][(kt,KF-1-kf,co)] (data gradient), bf16
__global__ void conv_window_weights_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wt, __nv_bfloat16* __restrict__ wd,
                                           int KT, int KF) {
  const int K = KT * KF * CW_C;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < K * CW_C; i += gridDim.x * blockDim.x) {
    const int co = i % CW_C;
    const int ci = (i / CW_C) % CW_C;
    const int kf = (i / (CW_C * CW_C)) % KF;
    const int kt = i / (CW_C * CW_C * KF);
    const __nv_bfloat16 v = __float2bfloat16(w[i]);
    wt[(size_t)co * K + (kt * KF + kf) * CW_C + ci] = v;
    wd[(size_t)ci * K + (kt * KF + (KF - 1 - kf)) * CW_C + co] = v;
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiledW)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// window view of a channel-last bf16 tensor [R][P][32] with row pitch `pitch` elements: {64, P, R} / {64 B, pitch}, box =
// PB positions x RB rows
static int make_tmap_window(CUtensorMap* out, const void* base, uint64_t P, uint64_t R, uint64_t pitch_elems, int PB, int RB) {
  static PFN_encodeTiledW enc = nullptr;
  if (!enc) {
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      enc = reinterpret_cast<PFN_encodeTiledW>(fp);
  }
  DS2_REQUIRE(enc != nullptr, DS2_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available from the driver");
  DS2_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (pitch_elems * 2) % 16 == 0, DS2_ERR_BAD_ARG,
              "conv window operand must be 16-byte aligned with a 16-byte multiple row pitch");
  cuuint64_t dims[3] = {64, P, R};
  cuuint64_t strides[2] = {CW_C * 2, pitch_elems * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)PB, (cuuint32_t)RB};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DS2_REQUIRE(r == CUDA_SUCCESS, DS2_ERR_CUDA, "cuTensorMapEncodeTiled (window view) failed with CUresult %d (P=%llu R=%llu pitch=%llu)",
              (int)r, (unsigned long long)P, (unsigned long long)R, (unsigned long long)pitch_elems);
  return DS2_OK;
}

static int window_shape_ok(const char* who, int B, int Ti, int Fi, int Ci, int Co, int KT, int KF, int st, int sf) {
  DS2_REQUIRE(B > 0 && Ti >= KT && Fi >= KF && KT > 0 && KF > 0 && st > 0, DS2_ERR_BAD_SHAPE, "%s: bad shape", who);
  DS2_REQUIRE(Ci == CW_C && Co == CW_C && sf == 1, DS2_ERR_UNSUPPORTED, "%s: needs 32 channels in and out and unit stride along "
              "frequency (Ci=%d Co=%d sf=%d)", who, Ci, Co, sf);
  DS2_REQUIRE(Fi <= 128, DS2_ERR_UNSUPPORTED, "%s: at most 128 positions per row (Fi=%d)", who, Fi);
  DS2_REQUIRE(KT * ((KF + 1) / 2) <= 32, DS2_ERR_UNSUPPORTED, "%s: filter %dx%d too large for the resident weight gradient", who, KT, KF);
  return DS2_OK;
}

static void fill_common(WinParams& p, int KT, int KF, int st) {
  p.KT = KT; p.st = st; p.nfull = KF / 2; p.half = KF & 1; p.kstride = KF * CW_C;
}

// positions per box, rows per box and position passes for rows of Pout positions: the split that fills most of the 128 MMA
// rows of a tile.  max_rows bounds RB (1: no packing).
static void choose_pack(int Pout, int max_rows, int* PB, int* RB, int* np) {
  double best = -1.0;
  for (int n = 1; n <= 16; ++n) {
    const int pb = (Pout + n - 1) / n;
    if (pb > 128) continue;
    int rb = 128 / pb;
    if (rb > max_rows) rb = max_rows;
    const double eff = (double)Pout * rb / (128.0 * n);
    if (eff > best + 1e-9) { best = eff; *PB = pb; *RB = rb; *np = n; }
  }
}
static constexpr int CW_MAX_RB = 16;
struct Pack {
  int PBf, RBf, npf;      // forward / weight gradient (rows of Fo positions, source rows st apart)
  int PBd, RBd, npd;      // data gradient (rows of Fi positions)
  int ptop, Rdy;          // padded dy: zero rows above an item's To rows, rows per item
};
static Pack pack_for(int Ti, int Fi, int KT, int KF, int st) {
  Pack k;
  const int To = (Ti - KT) / st + 1, Fo = Fi - KF + 1;
  choose_pack(Fo, st <= CW_MAXST ? CW_MAX_RB : 1, &k.PBf, &k.RBf, &k.npf);
  choose_pack(Fi, CW_MAX_RB, &k.PBd, &k.RBd, &k.npd);
  k.ptop = (KT - 1 + st - 1) / st;
  k.Rdy = k.ptop + To + k.ptop + 1 + (k.RBf > k.RBd ? k.RBf : k.RBd);
  return k;
}
// the source maps of the forward / weight gradient: one per residue of the row modulo st (packed rows), or the plain one
static int make_src_maps(WinMaps* m, int* nmaps, const void* a, int Fi, uint64_t rows, int st, int PB, int RB) {
  const uint64_t pitch = (uint64_t)Fi * CW_C;
  if (RB == 1) {
    *nmaps = 1;
    return make_tmap_window(&m->m[0], a, Fi, rows, pitch, PB, 1);
  }
  *nmaps = st;
  for (int r = 0; r < st; ++r) {
    if ((uint64_t)r >= rows) { m->m[r] = m->m[0]; continue; }
    int rc = make_tmap_window(&m->m[r], static_cast<const __nv_bfloat16*>(a) + (size_t)r * pitch, Fi, (rows - r + st - 1) / st, pitch * st, PB, RB);
    if (rc) return rc;
  }
  return DS2_OK;
}

}  // namespace ds2

using namespace ds2;

extern "C" int ds2_conv_window_supported(int Ci, int Co, int Fi, int KT, int KF, int sf) {
  return Ci == CW_C && Co == CW_C && sf == 1 && Fi <= 128 && Fi >= KF && KT * ((KF + 1) / 2) <= 32;
}

extern "C" int ds2_conv_window_weights(const float* w, void* wt, void* wd, int KT, int KF, void* stream) {
  DS2_REQUIRE(w && wt && wd && KT > 0 && KF > 0, DS2_ERR_BAD_ARG, "ds2_conv_window_weights: bad argument");
  conv_window_weights_kernel<<<(KT * KF * CW_C * CW_C + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w, static_cast<__nv_bfloat16*>(wt), static_cast<__nv_bfloat16*>(wd), KT, KF);
  DS2_LAUNCH_CHECK("conv_window_weights_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_conv_window_dy_layout(int Ti, int Fi, int KT, int KF, int st, int* ptop, int* rows_per_item) {
  DS2_REQUIRE(Ti >= KT && Fi >= KF && KT > 0 && KF > 0 && st > 0 && ptop && rows_per_item, DS2_ERR_BAD_ARG, "ds2_conv_window_dy_layout: bad argument");
  const Pack k = pack_for(Ti, Fi, KT, KF, st);
  *ptop = k.ptop;
  *rows_per_item = k.Rdy;
  return DS2_OK;
}

extern "C" int ds2_conv_window_pad_cast(const float* dy, void* out, int B, int Ti, int Fi, int KT, int KF, int st, void* stream) {
  DS2_REQUIRE(dy && out && B > 0 && Ti >= KT && Fi >= KF && KT > 0 && KF > 0 && st > 0, DS2_ERR_BAD_ARG, "ds2_conv_window_pad_cast: bad argument");
  const Pack k = pack_for(Ti, Fi, KT, KF, st);
  const int To = (Ti - KT) / st + 1, Fo = Fi - KF + 1, pad = KF - 1;
  long n4 = (long)B * k.Rdy * (Fo + 2 * pad) * (CW_C / 4);
  long grid = (n4 + 255) / 256;
  if (grid > (long)num_sms() * 16) grid = (long)num_sms() * 16;
  conv_pad_cast_kernel<<<(int)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, static_cast<__nv_bfloat16*>(out), B, To, Fo, pad, k.ptop, k.Rdy);
  DS2_LAUNCH_CHECK("conv_pad_cast_kernel");
  count_launch();
  return DS2_OK;
}

static int window_attr() {
  static unsigned long long done_devs = 0;
  if (first_use_on_device(&done_devs)) {
    DS2_CUDA(cudaFuncSetAttribute(conv_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CW_SMEM));
    DS2_CUDA(cudaFuncSetAttribute(conv_window_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CWG_SMEM));
  }
  return DS2_OK;
}

extern "C" int ds2_conv_window_fwd(const void* a, const void* wt, const float* bias, float* y, double* sums, int B, int Ti, int Fi, int KT,
                                   int KF, int st, void* stream) {
  DS2_REQUIRE(a && wt && y, DS2_ERR_BAD_ARG, "ds2_conv_window_fwd: null pointer");
  int rc = window_shape_ok("ds2_conv_window_fwd", B, Ti, Fi, CW_C, CW_C, KT, KF, st, 1);
  if (rc) return rc;
  if ((rc = window_attr())) return rc;
  const int To = (Ti - KT) / st + 1, Fo = Fi - KF + 1;
  const Pack k = pack_for(Ti, Fi, KT, KF, st);
  WinMaps ms = {};
  CUtensorMap tw;
  WinParams p = {};
  if ((rc = make_src_maps(&ms, &p.nmaps, a, Fi, (uint64_t)B * Ti, st, k.PBf, k.RBf))) return rc;
  if ((rc = make_tmap_2d_bf16(&tw, wt, (uint64_t)KT * KF * CW_C, CW_C, (uint64_t)KT * KF * CW_C, 64, CW_C))) return rc;
  fill_common(p, KT, KF, st);
  p.PB = k.PBf; p.RB = k.RBf; p.np = k.npf; p.G = (To + k.RBf - 1) / k.RBf; p.ncls = 1;
  p.out = y; p.bias = bias; p.n_tiles = B * p.G * p.np; p.Rout = To; p.Rsrc = Ti; p.Pout = Fo; p.win0 = 0; p.dgrad = 0;
  p.sums = sums;
  if (sums) DS2_CUDA(cudaMemsetAsync(sums, 0, 2 * CW_C * sizeof(double), static_cast<cudaStream_t>(stream)));
  const int grid = p.n_tiles < num_sms() ? p.n_tiles : num_sms();
  conv_window_kernel<<<grid, CW_THREADS, CW_SMEM, static_cast<cudaStream_t>(stream)>>>(ms, tw, p);
  DS2_LAUNCH_CHECK("conv_window_kernel<fwd>");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_conv_window_dgrad(const void* dyp, const void* wd, float* da, int B, int Ti, int Fi, int KT, int KF, int st,
                                     void* stream) {
  DS2_REQUIRE(dyp && wd && da, DS2_ERR_BAD_ARG, "ds2_conv_window_dgrad: null pointer");
  int rc = window_shape_ok("ds2_conv_window_dgrad", B, Ti, Fi, CW_C, CW_C, KT, KF, st, 1);
  if (rc) return rc;
  if ((rc = window_attr())) return rc;
  const int Fo = Fi - KF + 1, Pp = Fo + 2 * (KF - 1);
  const Pack k = pack_for(Ti, Fi, KT, KF, st);
  WinMaps ms = {};
  CUtensorMap tw;
  if ((rc = make_tmap_window(&ms.m[0], dyp, Pp, (uint64_t)B * k.Rdy, (uint64_t)Pp * CW_C, k.PBd, k.RBd))) return rc;
  for (int i = 1; i < CW_MAXST; ++i) ms.m[i] = ms.m[0];
  if ((rc = make_tmap_2d_bf16(&tw, wd, (uint64_t)KT * KF * CW_C, CW_C, (uint64_t)KT * KF * CW_C, 64, CW_C))) return rc;
  WinParams p = {};
  fill_common(p, KT, KF, st);
  const int rows_per_class = (Ti + st - 1) / st;
  p.PB = k.PBd; p.RB = k.RBd; p.np = k.npd; p.G = (rows_per_class + k.RBd - 1) / k.RBd; p.ncls = st; p.nmaps = 1; p.ptop = k.ptop;
  p.out = da; p.bias = nullptr; p.n_tiles = B * st * p.G * p.np; p.Rout = Ti; p.Rsrc = k.Rdy; p.Pout = Fi; p.win0 = 0; p.dgrad = 1;
  const int grid = p.n_tiles < num_sms() ? p.n_tiles : num_sms();
  conv_window_kernel<<<grid, CW_THREADS, CW_SMEM, static_cast<cudaStream_t>(stream)>>>(ms, tw, p);
  DS2_LAUNCH_CHECK("conv_window_kernel<dgrad>");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_conv_window_wgrad(const void* a, const void* dyp, float* dw, int B, int Ti, int Fi, int KT, int KF, int st,
                                     void* stream) {
  DS2_REQUIRE(a && dyp && dw, DS2_ERR_BAD_ARG, "ds2_conv_window_wgrad: null pointer");
  int rc = window_shape_ok("ds2_conv_window_wgrad", B, Ti, Fi, CW_C, CW_C, KT, KF, st, 1);
  if (rc) return rc;
  if ((rc = window_attr())) return rc;
  const int To = (Ti - KT) / st + 1, Fo = Fi - KF + 1, Pp = Fo + 2 * (KF - 1);
  const Pack k = pack_for(Ti, Fi, KT, KF, st);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DS2_CUDA(cudaMemsetAsync(dw, 0, (size_t)KT * KF * CW_C * CW_C * sizeof(float), s));
  WinMaps ms = {};
  CUtensorMap td;
  WinParams p = {};
  if ((rc = make_src_maps(&ms, &p.nmaps, a, Fi, (uint64_t)B * Ti, st, k.PBf, k.RBf))) return rc;
  if ((rc = make_tmap_window(&td, dyp, Pp, (uint64_t)B * k.Rdy, (uint64_t)Pp * CW_C, k.PBf, k.RBf))) return rc;
  fill_common(p, KT, KF, st);
  p.PB = k.PBf; p.RB = k.RBf; p.np = k.npf; p.G = (To + k.RBf - 1) / k.RBf; p.ncls = 1; p.ptop = k.ptop; p.Rdy = k.Rdy;
  p.out = dw; p.bias = nullptr; p.n_tiles = B * p.G * p.np; p.Rout = To; p.Rsrc = Ti; p.Pout = Fo; p.win0 = KF - 1; p.dgrad = 0;
  const int grid = p.n_tiles < num_sms() ? p.n_tiles : num_sms();
  conv_window_wgrad_kernel<<<grid, CW_THREADS, CWG_SMEM, s>>>(ms, td, p);
  DS2_LAUNCH_CHECK("conv_window_wgrad_kernel");
  count_launch();
  return DS2_OK;
}
