// Fused implicit-GEMM convolution of a ONE-channel input into 32 feature maps (conv1 of the reference: tf.nn.conv2d at
// src/deepSpeech_NCHW.py:110-119 / src/deepSpeech.py:114-121, filter [20,5,1,32], strides (2,2), followed by batch norm at
// :121-134): no patch matrix in global memory, and the per-channel sums the batch norm needs come out of the epilogue.
//
//   forward : y[b,t,f,co] = bias[co] + sum_{kt,kf} x[b, t*st+kt, f*sf+kf] w[kt,kf,0,co]
//             tile = 128 consecutive (t,f) positions of one batch item (flattened, so every MMA row is a real output and
//             the tile's y rows are one contiguous 16 KB block).  The CTA stages the input rows the tile touches in
//             shared memory, its threads write the patch rows as the K-major 128B-swizzled A operand (one 16-byte chunk
//             = the KF taps of one filter row, zero padded to 8: K = 8 KT), tcgen05.mma multiplies by the resident
//             filters [32 x 8 KT], the epilogue adds the bias, stores fp32 and keeps running per-channel sums of y and
//             y^2 in registers; one reduction + 64 double atomics per CTA at the end give the batch-norm statistics.
//   wgrad   : dW[kt,kf,0,co] = sum over positions of patch[pos][(kt,kf)] dy[pos][co], db[co] = sum dy[pos][co]:
//             the same patch rows read as the MN-major A operand, the tile's dy rows (rounded to bf16) as the MN-major B
//             operand, contraction over the 128 positions; the [8 KT x 32] partial sums of a CTA stay in TMEM until it has
//             run out of tiles.
// Several CTAs per SM hide each other's staging and MMA latencies (no intra-CTA pipeline).
#include "common.cuh"

namespace ds2 {
namespace c1 {

static constexpr int THREADS = 256;
static constexpr int CO = 32;
static constexpr int SLAB = 128 * 128;           // 128 rows x 64 bf16 of the A operand
static constexpr int WSLAB = CO * 128;           // 32 filters x 64 bf16
static constexpr int MAX_KT = 32;                // 4 slabs of 8 filter rows
static constexpr int MAX_KF = 8;

struct Params {
  const float* x;        // [B, Ti, Fi]
  const float* w;        // HWIO [KT, KF, 1, 32]
  const float* bias;     // [32] or null
  float* y;              // fwd: [B, To, Fo, 32]
  const float* dy;       // wgrad: [B, To, Fo, 32]
  float* dw;             // wgrad: HWIO [KT, KF, 1, 32], zeroed by the host
  float* db;             // wgrad, optional: [32], zeroed by the host
  double* sums;          // fwd, optional: [64] per-channel sum and sum of squares of y, zeroed by the host
  int B, Ti, Fi, To, Fo, KT, KF, st, sf;
  int tiles_per_item, n_tiles;
  int stage_rows, stage_pitch;   // staged input rows per tile (upper bound) and their pitch in floats
  int nslab;                     // 64-column slabs of the A operand: ceil(8 KT / 64)
};

struct Smem {
  uint32_t a_off, w_off, stage_off, dy_off, red_off, bar_off, total;
};
__host__ __device__ inline Smem smem_layout(int nslab_alloc, int stage_floats, bool wgrad) {
  Smem s;
  s.a_off = 0;
  s.w_off = s.a_off + nslab_alloc * SLAB;                        // fwd: filters; wgrad: unused
  s.dy_off = s.w_off + (wgrad ? 0 : nslab_alloc * WSLAB);        // wgrad: the tile's dy rows [128 x 128 B]
  s.stage_off = s.dy_off + (wgrad ? SLAB : 0);
  s.red_off = s.stage_off + 2 * ((stage_floats * 4 + 127) & ~127);     // two buffers: the next tile's rows load during this tile
  s.bar_off = s.red_off + 8 * 64 * 4;                            // per-warp partial channel sums
  s.total = s.bar_off + 64;
  return s;
}

// the input rows tile `tile` touches -> shared memory, asynchronously (4-byte cp.async: rows of Fi floats are not 16-byte
// aligned); the caller waits with stage_wait() one tile later
__device__ __forceinline__ void stage_rows_async(const Params& p, uint32_t stage_addr, int tile) {
  if (tile < p.n_tiles) {
    const int b = tile / p.tiles_per_item, pos0 = (tile - b * p.tiles_per_item) * 128;
    const int t0 = pos0 / p.Fo;
    const int last = min(pos0 + 127, p.To * p.Fo - 1);
    const int nrows = (last / p.Fo - t0) * p.st + p.KT;
    const float* src = p.x + ((size_t)b * p.Ti + (size_t)t0 * p.st) * p.Fi;
    for (int r = threadIdx.x >> 5; r < nrows; r += THREADS / 32) {
      const float* s = src + (size_t)r * p.Fi;
      const uint32_t d = stage_addr + r * p.stage_pitch * 4;
      for (int c = threadIdx.x & 31; c < p.Fi; c += 32)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + c * 4), "l"(s + c) : "memory");
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void stage_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// patch rows of the tile as the 128B-swizzled K-major operand: row m, 16-byte chunk kt = the KF taps of filter row kt.
// KFT > 0: the tap count at compile time (unrolled shared-space loads); 0: p.KF at run time.
template <int KFT>
__device__ __forceinline__ void build_patches(const Params& p, uint32_t stage_addr, uint32_t sA, int pos0, int t0, int npos_item) {
  const int m = threadIdx.x & 127;
  const int pos = pos0 + m;
  const bool valid = pos < npos_item;
  const int t = valid ? pos / p.Fo : t0, f = valid ? pos - t * p.Fo : 0;
  const uint32_t base = stage_addr + (((t - t0) * p.st) * p.stage_pitch + f * p.sf) * 4;
  const uint32_t row_addr = sA + m * 128;
  const uint32_t pitch_b = p.stage_pitch * 4;
  const int KF = KFT > 0 ? KFT : p.KF;
  const bool wide = KFT == 5 && (p.sf & 1) == 0;      // taps start at an even column: 8-byte aligned, two 64-bit loads + one
  for (int kt = threadIdx.x >> 7; kt < p.KT; kt += THREADS / 128) {
    float v[MAX_KF];
    if (wide) {
      const uint32_t a = base + kt * pitch_b;
      asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v[0]), "=f"(v[1]) : "r"(a) : "memory");
      asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v[2]), "=f"(v[3]) : "r"(a + 8) : "memory");
      v[4] = lds_f32(a + 16);
      v[5] = v[6] = v[7] = 0.f;
    } else {
#pragma unroll
      for (int k = 0; k < MAX_KF; ++k) v[k] = (k < KF) ? lds_f32(base + kt * pitch_b + k * 4) : 0.f;
    }
    uint4 pk;
    __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
    __nv_bfloat162 h2 = __floats2bfloat162_rn(v[4], v[5]), h3 = __floats2bfloat162_rn(v[6], v[7]);
    pk.x = *reinterpret_cast<uint32_t*>(&h0);
    pk.y = *reinterpret_cast<uint32_t*>(&h1);
    pk.z = *reinterpret_cast<uint32_t*>(&h2);
    pk.w = *reinterpret_cast<uint32_t*>(&h3);
    if (!valid) pk = make_uint4(0u, 0u, 0u, 0u);
    sts_v4(row_addr + (kt >> 3) * SLAB + (((kt & 7) ^ (m & 7)) << 4), pk);
  }
}

template <int KFT>
__global__ void __launch_bounds__(THREADS, 2)
conv1_fwd_kernel(Params p) {
  extern __shared__ __align__(1024) uint8_t c1_smem_raw[];
  const uint32_t sbase = (smem_u32(c1_smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = c1_smem_raw + (sbase - smem_u32(c1_smem_raw));
  const Smem L = smem_layout(p.nslab, p.stage_rows * p.stage_pitch, false);
  const uint32_t sA = sbase + L.a_off, sW = sbase + L.w_off;
  const uint32_t stage_bytes = (uint32_t)((p.stage_rows * p.stage_pitch * 4 + 127) & ~127);
  float* red = reinterpret_cast<float*>(gbase + L.red_off);
  const uint32_t mma_bar = sbase + L.bar_off;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gbase + L.bar_off + 16);
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(reinterpret_cast<uint64_t*>(gbase + L.bar_off), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 64);
  stage_rows_async(p, sbase + L.stage_off, blockIdx.x);
  // zero the operand buffers once (chunks of filter rows beyond KT are part of the contraction), then the filters:
  // row co, chunk kt = w[kt][0..KF-1][co]
  for (uint32_t o = threadIdx.x * 16u; o < (uint32_t)(p.nslab * (SLAB + WSLAB)); o += THREADS * 16u) sts_v4(sbase + o, make_uint4(0u, 0u, 0u, 0u));
  __syncthreads();
  for (int i = threadIdx.x; i < p.KT * CO; i += THREADS) {
    const int kt = i / CO, co = i - kt * CO;
    uint32_t h[4] = {0u, 0u, 0u, 0u};
    for (int k = 0; k < p.KF; ++k) {
      const __nv_bfloat16 b = __float2bfloat16(p.w[((size_t)kt * p.KF + k) * CO + co]);
      h[k >> 1] |= (uint32_t)(*reinterpret_cast<const unsigned short*>(&b)) << ((k & 1) * 16);
    }
    sts_v4(sW + (kt >> 3) * WSLAB + co * 128 + (((kt & 7) ^ (co & 7)) << 4), make_uint4(h[0], h[1], h[2], h[3]));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int q = warp & 3, half = warp >> 2;           // TMEM lane quarter and 16-channel half of this warp's epilogue
  float bias16[16], s1[16], s2[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    bias16[j] = p.bias ? __ldg(p.bias + half * 16 + j) : 0.f;
    s1[j] = 0.f;
    s2[j] = 0.f;
  }
  const int npos_item = p.To * p.Fo;
  const int ksteps = (p.KT + 1) / 2;                  // 16 K columns = two filter rows per MMA
  const uint32_t idesc = umma_idesc_bf16(128, CO, 0, 0);
  // the epilogue of one tile: accumulators (+ bias) -> y, running channel sums
  auto epilogue = [&](int tile, int acc) {
    const int b = tile / p.tiles_per_item, pos0 = (tile - b * p.tiles_per_item) * 128;
    uint32_t v[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(tmem_base + acc * CO + half * 16 + (static_cast<uint32_t>(q * 32) << 16))
        : "memory");
    tmem_ld_wait();
    tc_fence_before();
    const int pos = pos0 + q * 32 + lane;
    if (pos < npos_item) {
      float o[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        o[j] = __uint_as_float(v[j]) + bias16[j];
        s1[j] += o[j];
        s2[j] = fmaf(o[j], o[j], s2[j]);
      }
      float4* dst = reinterpret_cast<float4*>(p.y + ((size_t)b * npos_item + pos) * CO + half * 16);
#pragma unroll
      for (int j = 0; j < 4; ++j) dst[j] = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
    }
  };
  // Per tile: wait for its staged rows, start the next tile's, build the patch rows, issue the MMAs into one of two
  // accumulators, and while they run take the PREVIOUS tile's accumulator through the epilogue.
  uint32_t ph = 0;
  int buf = 0, it = 0, prev_tile = -1;
  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, buf ^= 1, ++it) {
    const int b = tile / p.tiles_per_item, pos0 = (tile - b * p.tiles_per_item) * 128;
    const int t0 = pos0 / p.Fo;
    (void)b;
    stage_wait();
    if (prev_tile >= 0) {              // the previous tile's MMAs have read the patch buffer (and filled its accumulator)
      mbar_wait(mma_bar, ph, 500);
      ph ^= 1;
      tc_fence_after();
    }
    __syncthreads();
    stage_rows_async(p, sbase + L.stage_off + (buf ^ 1) * stage_bytes, tile + gridDim.x);
    build_patches<KFT>(p, sbase + L.stage_off + buf * stage_bytes, sA, pos0, t0, npos_item);
    fence_proxy_async();
    __syncthreads();
    if (warp == 0 && elect_one_sync()) {
      tc_fence_after();
      for (int ks = 0; ks < ksteps; ++ks)
        umma_bf16(tmem_base + (it & 1) * CO, umma_smem_desc_sw128(sA + (ks >> 2) * SLAB + (ks & 3) * 32, 16, 1024),
                  umma_smem_desc_sw128(sW + (ks >> 2) * WSLAB + (ks & 3) * 32, 16, 1024), idesc, ks > 0 ? 1u : 0u);
      umma_commit(mma_bar);
    }
    __syncwarp();
    if (prev_tile >= 0) epilogue(prev_tile, (it - 1) & 1);
    prev_tile = tile;
  }
  if (prev_tile >= 0) {
    mbar_wait(mma_bar, ph, 501);
    tc_fence_after();
    epilogue(prev_tile, (it - 1) & 1);
  }
  if (p.sums) {
    // per-channel sums: lanes -> warp, the four quarter warps of a half -> one double atomic per channel and moment
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      s1[j] = warp_sum(s1[j]);
      s2[j] = warp_sum(s2[j]);
    }
    if (lane == 0) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        red[warp * 64 + j] = s1[j];
        red[warp * 64 + 32 + j] = s2[j];
      }
    }
    __syncthreads();
    if (threadIdx.x < 64) {
      const int mom = threadIdx.x >> 5, c = threadIdx.x & 31, h = c >> 4, j = c & 15;
      double a = 0.0;
      for (int w4 = 0; w4 < 4; ++w4) a += (double)red[(h * 4 + w4) * 64 + mom * 32 + j];
      atomicAdd(p.sums + mom * CO + c, a);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

// ------------------------------------------------------------------ weight and bias gradient
template <int KFT>
__global__ void __launch_bounds__(THREADS, 2)
conv1_wgrad_kernel(Params p) {
  extern __shared__ __align__(1024) uint8_t c1_smem_raw[];
  const uint32_t sbase = (smem_u32(c1_smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = c1_smem_raw + (sbase - smem_u32(c1_smem_raw));
  const int nblk = (p.nslab + 1) / 2;                 // MMA blocks of two 64-wide slabs (M = 128 window elements)
  // with an odd slab count the last block's second half reads whatever follows the last slab (the dy tile): rows 64..127
  // of that block's accumulators are never used
  const Smem L = smem_layout(p.nslab, p.stage_rows * p.stage_pitch, true);
  const uint32_t sA = sbase + L.a_off, sDy = sbase + L.dy_off;
  const uint32_t stage_bytes = (uint32_t)((p.stage_rows * p.stage_pitch * 4 + 127) & ~127);
  float* red = reinterpret_cast<float*>(gbase + L.red_off);
  const uint32_t mma_bar = sbase + L.bar_off;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gbase + L.bar_off + 16);
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(reinterpret_cast<uint64_t*>(gbase + L.bar_off), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 64);
  stage_rows_async(p, sbase + L.stage_off, blockIdx.x);
  for (uint32_t o = threadIdx.x * 16u; o < (uint32_t)(p.nslab * SLAB + SLAB); o += THREADS * 16u) sts_v4(sbase + o, make_uint4(0u, 0u, 0u, 0u));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int m = threadIdx.x & 127, half = threadIdx.x >> 7;   // dy row and 16-channel half this thread converts
  float dbs[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) dbs[j] = 0.f;
  const int npos_item = p.To * p.Fo;
  const uint32_t idesc = umma_idesc_bf16(128, CO, 1, 1);
  uint32_t ph = 0;
  bool first = true;
  const bool has_tiles = (int)blockIdx.x < p.n_tiles;
  int buf = 0;
  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, buf ^= 1) {
    const int b = tile / p.tiles_per_item, pos0 = (tile - b * p.tiles_per_item) * 128;
    const int t0 = pos0 / p.Fo;
    // this thread's half row of dy (issued before the wait: the loads overlap the previous tile's MMAs)
    const int pos = pos0 + m;
    float4 g[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) g[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (pos < npos_item) {
      const float4* src = reinterpret_cast<const float4*>(p.dy + ((size_t)b * npos_item + pos) * CO + half * 16);
#pragma unroll
      for (int j = 0; j < 4; ++j) g[j] = __ldg(src + j);
    }
    if (!first) {                                     // the previous tile's MMAs have read the operand buffers
      mbar_wait(mma_bar, ph, 510);
      ph ^= 1;
    }
    stage_wait();
    __syncthreads();
    stage_rows_async(p, sbase + L.stage_off + (buf ^ 1) * stage_bytes, tile + gridDim.x);
    build_patches<KFT>(p, sbase + L.stage_off + buf * stage_bytes, sA, pos0, t0, npos_item);
    {
      const float* gf = reinterpret_cast<const float*>(g);
      uint32_t h[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dbs[2 * j] += gf[2 * j];
        dbs[2 * j + 1] += gf[2 * j + 1];
        __nv_bfloat162 t2 = __floats2bfloat162_rn(gf[2 * j], gf[2 * j + 1]);
        h[j] = *reinterpret_cast<uint32_t*>(&t2);
      }
      sts_v4(sDy + m * 128 + (((half * 2) ^ (m & 7)) << 4), make_uint4(h[0], h[1], h[2], h[3]));
      sts_v4(sDy + m * 128 + (((half * 2 + 1) ^ (m & 7)) << 4), make_uint4(h[4], h[5], h[6], h[7]));
    }
    fence_proxy_async();
    __syncthreads();
    if (warp == 0 && elect_one_sync()) {
      tc_fence_after();
      for (int blk = 0; blk < nblk; ++blk)
        for (int k = 0; k < 8; ++k)       // contraction over the tile's 128 positions, 16 (16 x 128 B) per step
          umma_bf16(tmem_base + blk * CO, umma_smem_desc_sw128(sA + blk * 2 * SLAB + k * 2048, SLAB, 1024),
                    umma_smem_desc_sw128(sDy + k * 2048, SLAB, 1024), idesc, (first && k == 0) ? 0u : 1u);
      umma_commit(mma_bar);
    }
    __syncwarp();
    first = false;
  }
  if (has_tiles) {
    mbar_wait(mma_bar, ph, 511);
    tc_fence_after();
    const int q = warp & 3, ch = warp >> 2;
    for (int blk = 0; blk < nblk; ++blk) {
      uint32_t v[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
            "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
          : "r"(tmem_base + blk * CO + ch * 16 + (static_cast<uint32_t>(q * 32) << 16))
          : "memory");
      tmem_ld_wait();
      const int k = blk * 128 + q * 32 + lane;        // window element: filter row k / 8, tap k % 8
      const int kt = k >> 3, kf = k & 7;
      if (kt < p.KT && kf < p.KF) {
        float* dst = p.dw + ((size_t)kt * p.KF + kf) * CO + ch * 16;
#pragma unroll
        for (int j = 0; j < 16; ++j) atomicAdd(dst + j, __uint_as_float(v[j]));
      }
    }
    if (p.db) {
#pragma unroll
      for (int j = 0; j < 16; ++j) dbs[j] = warp_sum(dbs[j]);
      if (lane == 0) {
#pragma unroll
        for (int j = 0; j < 16; ++j) red[warp * 16 + j] = dbs[j];
      }
    }
  }
  __syncthreads();
  if (p.db && has_tiles && threadIdx.x < CO) {
    // warps 0..3 hold rows 0..127 of channel half 0, warps 4..7 of half 1
    const int h = threadIdx.x >> 4, j = threadIdx.x & 15;
    float a = 0.f;
    for (int w4 = 0; w4 < 4; ++w4) a += red[(h * 4 + w4) * 16 + j];
    atomicAdd(p.db + threadIdx.x, a);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

}  // namespace c1
}  // namespace ds2

// ------------------------------------------------------------------ host side
using namespace ds2;
using namespace ds2::c1;

static int conv1_geometry(int Ti, int Fi, int KT, int KF, int st, int sf, Params* p, size_t* smem_fwd, size_t* smem_wgrad) {
  if (KT <= 0 || KF <= 0 || KT > MAX_KT || KF > MAX_KF || st <= 0 || sf <= 0 || Ti < KT || Fi < KF) return 0;
  const int To = (Ti - KT) / st + 1, Fo = (Fi - KF) / sf + 1;
  const int nslab = (KT + 7) / 8;
  const int nt_max = (127 + Fo - 1) / Fo + 1;                 // output rows a 128-position tile can touch
  const int rows = (nt_max - 1) * st + KT;
  const int pitch = (Fi + 3) & ~3;
  const size_t sf_bytes = smem_layout(nslab, rows * pitch, false).total + 1024;
  const size_t sw_bytes = smem_layout(nslab, rows * pitch, true).total + 1024;
  if (sf_bytes > 112 * 1024 || sw_bytes > 112 * 1024) return 0;  // two CTAs per SM
  if (p) {
    p->To = To; p->Fo = Fo; p->nslab = nslab; p->stage_rows = rows; p->stage_pitch = pitch;
    p->tiles_per_item = (To * Fo + 127) / 128;
  }
  if (smem_fwd) *smem_fwd = sf_bytes;
  if (smem_wgrad) *smem_wgrad = sw_bytes;
  return 1;
}

extern "C" int ds2_conv1_fused_supported(int Ci, int Co, int Ti, int Fi, int KT, int KF, int st, int sf) {
  return Ci == 1 && Co == CO && conv1_geometry(Ti, Fi, KT, KF, st, sf, nullptr, nullptr, nullptr);
}

static int conv1_attr() {
  static unsigned long long done_devs = 0;
  if (first_use_on_device(&done_devs)) {
    DS2_CUDA(cudaFuncSetAttribute(conv1_fwd_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    DS2_CUDA(cudaFuncSetAttribute(conv1_fwd_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    DS2_CUDA(cudaFuncSetAttribute(conv1_wgrad_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    DS2_CUDA(cudaFuncSetAttribute(conv1_wgrad_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
  }
  return DS2_OK;
}

extern "C" int ds2_conv1_fused_fwd(const float* x, const float* w, const float* bias, float* y, double* sums, int B, int Ti, int Fi, int KT,
                                   int KF, int st, int sf, void* stream) {
  DS2_REQUIRE(x && w && y && B > 0, DS2_ERR_BAD_ARG, "ds2_conv1_fused_fwd: bad argument");
  Params p = {};
  size_t smem = 0;
  DS2_REQUIRE(conv1_geometry(Ti, Fi, KT, KF, st, sf, &p, &smem, nullptr), DS2_ERR_UNSUPPORTED,
              "ds2_conv1_fused_fwd: shape not taken (Ti=%d Fi=%d filter %dx%d strides %d,%d)", Ti, Fi, KT, KF, st, sf);
  int rc = conv1_attr();
  if (rc) return rc;
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (sums) DS2_CUDA(cudaMemsetAsync(sums, 0, 2 * CO * sizeof(double), cs));
  p.x = x; p.w = w; p.bias = bias; p.y = y; p.sums = sums;
  p.B = B; p.Ti = Ti; p.Fi = Fi; p.KT = KT; p.KF = KF; p.st = st; p.sf = sf;
  p.n_tiles = B * p.tiles_per_item;
  const int grid = p.n_tiles < 2 * num_sms() ? p.n_tiles : 2 * num_sms();
  if (KF == 5) conv1_fwd_kernel<5><<<grid, THREADS, smem, cs>>>(p);      // the reference's [20,5] filter
  else conv1_fwd_kernel<0><<<grid, THREADS, smem, cs>>>(p);
  DS2_LAUNCH_CHECK("conv1_fwd_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_conv1_fused_wgrad(const float* x, const float* dy, float* dw, float* db, int B, int Ti, int Fi, int KT, int KF, int st,
                                     int sf, void* stream) {
  DS2_REQUIRE(x && dy && dw && B > 0, DS2_ERR_BAD_ARG, "ds2_conv1_fused_wgrad: bad argument");
  Params p = {};
  size_t smem = 0;
  DS2_REQUIRE(conv1_geometry(Ti, Fi, KT, KF, st, sf, &p, nullptr, &smem), DS2_ERR_UNSUPPORTED,
              "ds2_conv1_fused_wgrad: shape not taken (Ti=%d Fi=%d filter %dx%d strides %d,%d)", Ti, Fi, KT, KF, st, sf);
  int rc = conv1_attr();
  if (rc) return rc;
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  DS2_CUDA(cudaMemsetAsync(dw, 0, (size_t)KT * KF * CO * sizeof(float), cs));
  if (db) DS2_CUDA(cudaMemsetAsync(db, 0, CO * sizeof(float), cs));
  p.x = x; p.dy = dy; p.dw = dw; p.db = db;
  p.B = B; p.Ti = Ti; p.Fi = Fi; p.KT = KT; p.KF = KF; p.st = st; p.sf = sf;
  p.n_tiles = B * p.tiles_per_item;
  const int grid = p.n_tiles < 2 * num_sms() ? p.n_tiles : 2 * num_sms();
  if (KF == 5) conv1_wgrad_kernel<5><<<grid, THREADS, smem, cs>>>(p);
  else conv1_wgrad_kernel<0><<<grid, THREADS, smem, cs>>>(p);
  DS2_LAUNCH_CHECK("conv1_wgrad_kernel");
  count_launch();
  return DS2_OK;
}
