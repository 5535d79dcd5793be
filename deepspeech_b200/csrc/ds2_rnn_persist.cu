// Persistent, weights-resident recurrent kernel (clipped-ReLU RNN), forward and backward.
//
// Replaces the tf.while_loop of tf.nn.bidirectional_dynamic_rnn around CustomRNNCell2
// (src/custom_ops.py:48-72, :91-93): per timestep  h_t = clip(relu(s*Wx_t + U h_{t-1} + b)).
// The per-step product U.h is a [H x H] x [H x B] contraction on the tensor cores:
//   * the H output units are tiled by 128 (tcgen05 M=128), the batch is the MMA N dimension;
//   * the reduction dimension is split over a thread-block cluster of 4 CTAs, so that each CTA keeps
//     its [128 x H/4] bf16 slice of U RESIDENT in shared memory for all T steps (loaded once by TMA);
//   * each step: TMA-load the CTA's K-slice of h_{t-1} -> tcgen05.mma into TMEM -> tcgen05.ld ->
//     partial sums exchanged through distributed shared memory (each CTA finishes 32 of the 128 rows)
//     -> epilogue (add s*Wx_t + b, clip, length mask) -> bf16 h_t to global;
//   * both directions run concurrently as independent groups of CTAs; timesteps are ordered by a
//     per-direction grid barrier (release/acquire counter in global memory).
// The backward kernel is the same machine with A = U^T (pre-transposed by the caller-side wrapper),
// B operand = dz_{t+1}, and the epilogue dz_t = [0 < h_t < cap] * (dout_t + U^T dz_{t+1}).
#include <stdlib.h>

#include "common.cuh"

namespace ds2 {

static constexpr int KS = 4;            // cluster size = K split
static constexpr int RNN_THREADS = 160; // warp 0: control (TMA + MMA issue); warps 1-4: epilogue

struct RnnPersistParams {
  const float* add_in;     // fwd: wx [T,B,ND*H] ; bwd: dout [T,B,H]
  const float* bias;       // fwd: [ND,H]
  const int* lens;         // [B]
  const __nv_bfloat16* hs_in;   // bwd: hs [ND,T,B,H] (activation mask)
  __nv_bfloat16* out;      // fwd: hs [ND,T,B,H] ; bwd: dz [T,B,ND*H]
  unsigned int* ctr;       // [ND] step counters, 32 uints apart
  long long* prof;         // optional per-step clock64 stamps of CTA (0,0,0) (DS2_RNN_PROF=1), 8 per step
  int T, B, H, ND, reverse_mask;
  int tiles_m, ksl, nbox;  // M tiles per direction, K-slice length (multiple of 64), boxes per slice
  float scale, cap;
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
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <bool BWD>
__global__ void __launch_bounds__(RNN_THREADS, 1)
rnn_persist_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b0,
                   const __grid_constant__ CUtensorMap tm_b1, RnnPersistParams p) {
  extern __shared__ uint8_t rnn_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(rnn_smem_raw) + 1023) & ~uintptr_t(1023));
  const int Bn = p.B;
  const int box_b_bytes = Bn * 128;                         // one [B x 64] bf16 box of the moving operand
  uint8_t* sA = smem;                                       // nbox x [128 x 64] bf16 (SW128, K-major)
  uint8_t* sB = sA + p.nbox * 16384;                        // nbox x [B x 64] bf16
  float* slab = reinterpret_cast<float*>(sB + p.nbox * box_b_bytes);   // [KS][B][32] fp32 partial sums
  uint64_t* bars = reinterpret_cast<uint64_t*>(slab + KS * Bn * 32);
  uint64_t* a_bar = bars;          // U slice landed
  uint64_t* h_bar = bars + 1;      // moving operand landed (per step)
  uint64_t* mma_bar = bars + 2;    // accumulators ready (per step)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                  // K slice owned by this CTA, rows it finishes
  const int mt = blockIdx.y, d = blockIdx.z;
  const int m0 = mt * 128;
  const int rev = (p.reverse_mask >> d) & 1;
  const int T = p.T, H = p.H, ND = p.ND;
  const CUtensorMap* tm_b = d ? &tm_b1 : &tm_b0;
  unsigned int* ctr = p.ctr + d * 32;
  const unsigned int ctas_per_dir = KS * p.tiles_m;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_a);
    tma_prefetch_desc(tm_b);
    mbar_init(a_bar, 1);
    mbar_init(h_bar, 1);
    mbar_init(mma_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // every CTA of the cluster must have initialised its barriers / be alive before DSMEM traffic
  cluster_arrive();
  cluster_wait();

  if (warp == 0) {
    // =========================== control warp: TMA + MMA issue ===========================
    if (lane == 0) {
      mbar_arrive_expect_tx(a_bar, p.nbox * 16384);
      for (int i = 0; i < p.nbox; ++i) tma_load_2d(sA + i * 16384, &tm_a, a_bar, rank * p.ksl + i * 64, d * H + m0);
      mbar_wait(a_bar, 0, 10);
    }
    __syncwarp();
    const uint32_t idesc = umma_idesc_bf16(128, Bn, 0, 0);
    uint32_t ph = 0;
    for (int step = 1; step < T; ++step) {
      const int t = BWD ? (rev ? step : T - 1 - step) : (rev ? T - 1 - step : step);
      const int tsrc = BWD ? (rev ? t - 1 : t + 1) : (rev ? t + 1 : t - 1);   // step whose result feeds this one
      if (lane == 0) {
        // wait until every CTA of this direction has published the previous step
        const unsigned int need = ctas_per_dir * (unsigned int)step;
        if (ld_acquire_u32(ctr) < need) {
          SpinGuard g;
          while (ld_acquire_u32(ctr) < need) g.check(20);
        }
        if (p.prof && blockIdx.x + blockIdx.y + blockIdx.z == 0 && step < 1024) p.prof[step * 8 + 0] = dev_clock();
        asm volatile("fence.proxy.async;" ::: "memory");      // generic-proxy acquire -> async-proxy (TMA) reads
        tc_fence_after();
        mbar_arrive_expect_tx(h_bar, p.nbox * box_b_bytes);
        for (int i = 0; i < p.nbox; ++i)
          tma_load_2d(sB + i * box_b_bytes, tm_b, h_bar, rank * p.ksl + i * 64, tsrc * Bn);
        mbar_wait(h_bar, ph, 30);
        if (p.prof && blockIdx.x + blockIdx.y + blockIdx.z == 0 && step < 1024) p.prof[step * 8 + 1] = dev_clock();
        tc_fence_after();
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
        for (int i = 0; i < p.nbox; ++i) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t adesc = umma_smem_desc_sw128(a0 + i * 16384 + k * 32, 16, 1024);
            const uint64_t bdesc = umma_smem_desc_sw128(b0 + i * box_b_bytes + k * 32, 16, 1024);
            umma_bf16(tmem_base, adesc, bdesc, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(mma_bar);
        if (p.prof && blockIdx.x + blockIdx.y + blockIdx.z == 0 && step < 1024) p.prof[step * 8 + 2] = dev_clock();
      }
      __syncwarp();
      ph ^= 1;
      cluster_arrive();        // matches the epilogue warps' exchange barrier of this step
      cluster_wait();
    }
  } else {
    // =========================== epilogue warps ===========================
    const int q = warp & 3;                      // TMEM lane quarter = rows [32q, 32q+32) of the tile
    const int et = threadIdx.x - 32;             // 0..127
    const int jq = et & 3;                       // 8-wide chunk of this CTA's 32 rows
    const int j = m0 + 32 * (int)rank + 8 * jq;  // first of 8 consecutive output units
    const bool jvalid = j < H;                   // H % 8 == 0 -> all-or-nothing
    const int nb = Bn / 32;                      // batch groups handled per thread (1 or 2)
    float bias8[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) bias8[i] = (!BWD && jvalid) ? p.bias[d * H + j + i] : 0.f;
    int len_b[2];
    for (int g = 0; g < nb; ++g) len_b[g] = p.lens[(et >> 2) + 32 * g];
    const uint32_t slab_local = smem_u32(slab);
    uint32_t ph = 0;
    for (int step = 0; step < T; ++step) {
      const int t = BWD ? (rev ? step : T - 1 - step) : (rev ? T - 1 - step : step);
      // prefetch the additive term (s*Wx_t or dout_t) and, for the backward, h_t for the mask
      float add[2][8];
      uint4 hm[2];
      for (int g = 0; g < nb; ++g) {
        const int b = (et >> 2) + 32 * g;
        if (jvalid) {
          const float* src = BWD ? p.add_in + ((size_t)t * Bn + b) * H + j
                                 : p.add_in + ((size_t)t * Bn + b) * ND * H + (size_t)d * H + j;
          const float4 v0 = __ldg(reinterpret_cast<const float4*>(src));
          const float4 v1 = __ldg(reinterpret_cast<const float4*>(src) + 1);
          add[g][0] = v0.x; add[g][1] = v0.y; add[g][2] = v0.z; add[g][3] = v0.w;
          add[g][4] = v1.x; add[g][5] = v1.y; add[g][6] = v1.z; add[g][7] = v1.w;
          if (BWD) hm[g] = *reinterpret_cast<const uint4*>(p.hs_in + (((size_t)d * T + t) * Bn + b) * H + j);
        }
      }
      float acc[2][8];
#pragma unroll
      for (int g = 0; g < 2; ++g)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[g][i] = 0.f;
      if (step > 0) {
        mbar_wait(mma_bar, ph, 40);
        const bool pr = p.prof && et == 0 && blockIdx.x + blockIdx.y + blockIdx.z == 0 && step < 1024;
        if (pr) p.prof[step * 8 + 3] = dev_clock();
        ph ^= 1;
        tc_fence_after();
        // scatter this warp's 32 rows x B partial sums to the CTA that finishes those rows (rank q)
        const uint32_t dst = map_to_rank(slab_local, (uint32_t)q);
        for (int g = 0; g < nb; ++g) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + g * 32 + (static_cast<uint32_t>(q * 32) << 16), r);
          tmem_ld_wait();
#pragma unroll
          for (int bb = 0; bb < 32; ++bb)
            st_cluster_f32(dst + ((rank * Bn + g * 32 + bb) * 32 + lane) * 4, __uint_as_float(r[bb]));
        }
        tc_fence_before();
        if (pr) p.prof[step * 8 + 4] = dev_clock();
        cluster_arrive();
        cluster_wait();
        if (pr) p.prof[step * 8 + 5] = dev_clock();
        // reduce the KS partial slabs for this thread's (b, 8 rows)
        for (int g = 0; g < nb; ++g) {
          const int b = (et >> 2) + 32 * g;
#pragma unroll
          for (int s = 0; s < KS; ++s) {
            const float4* sp = reinterpret_cast<const float4*>(slab + (s * Bn + b) * 32 + 8 * jq);
            const float4 v0 = sp[0], v1 = sp[1];
            acc[g][0] += v0.x; acc[g][1] += v0.y; acc[g][2] += v0.z; acc[g][3] += v0.w;
            acc[g][4] += v1.x; acc[g][5] += v1.y; acc[g][6] += v1.z; acc[g][7] += v1.w;
          }
        }
      }
      if (jvalid) {
        for (int g = 0; g < nb; ++g) {
          const int b = (et >> 2) + 32 * g;
          const bool live = t < len_b[g];
          float o[8];
          if (!BWD) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float z = fmaf(p.scale, add[g][i], acc[g][i] + bias8[i]);
              o[i] = live ? fminf(fmaxf(z, 0.f), p.cap) : 0.f;
            }
          } else {
            const __nv_bfloat16* hb = reinterpret_cast<const __nv_bfloat16*>(&hm[g]);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float h = __bfloat162float(hb[i]);
              o[i] = (live && h > 0.f && h < p.cap) ? (add[g][i] + acc[g][i]) : 0.f;
            }
          }
          uint4 pk;
          __nv_bfloat162 b0 = __floats2bfloat162_rn(o[0], o[1]), b1 = __floats2bfloat162_rn(o[2], o[3]);
          __nv_bfloat162 b2 = __floats2bfloat162_rn(o[4], o[5]), b3 = __floats2bfloat162_rn(o[6], o[7]);
          pk.x = *reinterpret_cast<uint32_t*>(&b0); pk.y = *reinterpret_cast<uint32_t*>(&b1);
          pk.z = *reinterpret_cast<uint32_t*>(&b2); pk.w = *reinterpret_cast<uint32_t*>(&b3);
          __nv_bfloat16* dstp = BWD ? p.out + ((size_t)t * Bn + b) * ND * H + (size_t)d * H + j
                                    : p.out + (((size_t)d * T + t) * Bn + b) * H + j;
          *reinterpret_cast<uint4*>(dstp) = pk;
        }
      }
      // publish: all 128 epilogue threads' stores -> one release increment of the direction counter
      const bool pr2 = p.prof && et == 0 && blockIdx.x + blockIdx.y + blockIdx.z == 0 && step < 1024;
      if (pr2) p.prof[step * 8 + 6] = dev_clock();
      // CTA barrier orders the 128 threads' h_t stores before thread 0's gpu-scope release (cumulativity)
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (et == 0) red_release_add(ctr, 1u);
      if (pr2) p.prof[step * 8 + 7] = dev_clock();
    }
  }
  // no CTA may exit while a peer can still write into its shared memory
  tc_fence_before();
  cluster_arrive();
  cluster_wait();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

// ------------------------------------------------------------------ host side
static int persist_geometry(int B, int H, int ND, int* tiles_m, int* ksl, int* nbox, size_t* smem) {
  if (!(B == 32 || B == 64) || H % 8 != 0 || H < 8 || (ND != 1 && ND != 2)) return 0;
  const int tm = (H + 127) / 128;
  const int slice = (H + KS - 1) / KS;
  const int l = ((slice + 63) / 64) * 64;
  const int nb = l / 64;
  const size_t bytes = (size_t)nb * 16384 + (size_t)nb * B * 128 + (size_t)KS * B * 32 * 4 + 64 + 1024;
  if (bytes > 225 * 1024) return 0;
  if (KS * tm * ND > 128) return 0;           // must be co-resident in clusters of 4 (148 SMs, 1 CTA / SM)
  if (tiles_m) *tiles_m = tm;
  if (ksl) *ksl = l;
  if (nbox) *nbox = nb;
  if (smem) *smem = bytes;
  return 1;
}

int rnn_persist_supported(int B, int H, int ND) { return persist_geometry(B, H, ND, nullptr, nullptr, nullptr, nullptr); }

size_t rnn_persist_workspace_bytes(int T, int B, int H, int ND) {
  (void)T; (void)B;
  // 1 KB of step counters (one 128-byte line per direction) + room for U^T (bf16) used by the backward
  return 1024 + 65536 + (size_t)ND * H * H * 2;
}

template <bool BWD>
static int persist_launch(const CUtensorMap& ta, const CUtensorMap& tb0, const CUtensorMap& tb1, RnnPersistParams& p, size_t smem,
                          cudaStream_t st) {
  auto kern = rnn_persist_kernel<BWD>;
  static bool attr_done[2] = {false, false};
  if (!attr_done[BWD]) {
    DS2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
    attr_done[BWD] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(KS, p.tiles_m, p.ND);
  cfg.blockDim = dim3(RNN_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = KS;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;       // all CTAs co-resident (they spin on each other)
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = getenv("DS2_RNN_NO_COOP") ? 1 : 2;   // profilers cannot replay cooperative launches
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb0, tb1, p);
  if (e != cudaSuccess) {
    // some driver versions reject cooperative + cluster launches: the grid (<= 128 CTAs, 1 per SM)
    // fits the device anyway, so fall back to a plain cluster launch
    (void)cudaGetLastError();
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, kern, ta, tb0, tb1, p);
  }
  if (e != cudaSuccess) return cuda_fail(e, BWD ? "rnn_persist_kernel<bwd>" : "rnn_persist_kernel<fwd>");
  count_launch();
  return DS2_OK;
}

int rnn_relu_fwd_persist(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs, int T, int B,
                         int H, int ND, int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st) {
  RnnPersistParams p = {};
  size_t smem;
  DS2_REQUIRE(persist_geometry(B, H, ND, &p.tiles_m, &p.ksl, &p.nbox, &smem), DS2_ERR_UNSUPPORTED, "rnn persist: unsupported shape");
  DS2_REQUIRE(ws && ws_bytes >= 1024 + 65536, DS2_ERR_WORKSPACE, "ds2_rnn_relu_fwd: workspace of %d bytes required", 1024 + 65536);
  DS2_CUDA(cudaMemsetAsync(ws, 0, 1024, st));
  CUtensorMap ta, tb0, tb1;
  int rc = make_tmap_2d_bf16(&ta, u, H, (uint64_t)ND * H, H, 64, 128);
  if (rc) return rc;
  const __nv_bfloat16* hsb = static_cast<const __nv_bfloat16*>(hs);
  rc = make_tmap_2d_bf16(&tb0, hsb, H, (uint64_t)T * B, H, 64, B);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tb1, hsb + (ND > 1 ? (size_t)T * B * H : 0), H, (uint64_t)T * B, H, 64, B);
  if (rc) return rc;
  p.add_in = wx; p.bias = bias; p.lens = lens; p.hs_in = nullptr; p.out = static_cast<__nv_bfloat16*>(hs);
  p.ctr = static_cast<unsigned int*>(ws);
  p.prof = getenv("DS2_RNN_PROF") ? reinterpret_cast<long long*>(static_cast<char*>(ws) + 1024) : nullptr;
  p.T = T; p.B = B; p.H = H; p.ND = ND; p.reverse_mask = reverse_mask; p.scale = scale; p.cap = cap;
  return persist_launch<false>(ta, tb0, tb1, p, smem, st);
}

// The backward contraction is U^T dz: U is transposed per direction into the workspace first
// (ut[d][k][j] = U[d][j][k], 2*ND*H*H bytes, one small kernel per direction), then used as the resident A operand.
int rnn_relu_bwd_persist(const float* dout, const void* u, const int* lens, const void* hs, void* dz, int T, int B, int H, int ND,
                         int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st) {
  RnnPersistParams p = {};
  size_t smem;
  DS2_REQUIRE(persist_geometry(B, H, ND, &p.tiles_m, &p.ksl, &p.nbox, &smem), DS2_ERR_UNSUPPORTED, "rnn persist: unsupported shape");
  const size_t need = rnn_persist_workspace_bytes(T, B, H, ND);
  DS2_REQUIRE(ws && ws_bytes >= need, DS2_ERR_WORKSPACE, "ds2_rnn_relu_bwd: workspace of %zu bytes required", need);
  DS2_CUDA(cudaMemsetAsync(ws, 0, 1024, st));
  __nv_bfloat16* ut = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(ws) + 1024 + 65536);
  int rc;
  for (int d = 0; d < ND; ++d) {
    rc = ds2_transpose(static_cast<const __nv_bfloat16*>(u) + (size_t)d * H * H, DS2_BF16, H, H, H, ut + (size_t)d * H * H, DS2_BF16, H, st);
    if (rc) return rc;
  }
  CUtensorMap ta, tb0, tb1;
  rc = make_tmap_2d_bf16(&ta, ut, H, (uint64_t)ND * H, H, 64, 128);
  if (rc) return rc;
  const __nv_bfloat16* dzb = static_cast<const __nv_bfloat16*>(dz);
  rc = make_tmap_2d_bf16(&tb0, dzb, H, (uint64_t)T * B, (uint64_t)ND * H, 64, B);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tb1, dzb + (ND > 1 ? H : 0), H, (uint64_t)T * B, (uint64_t)ND * H, 64, B);
  if (rc) return rc;
  p.add_in = dout; p.bias = nullptr; p.lens = lens; p.hs_in = static_cast<const __nv_bfloat16*>(hs);
  p.out = static_cast<__nv_bfloat16*>(dz);
  p.ctr = static_cast<unsigned int*>(ws);
  p.prof = getenv("DS2_RNN_PROF") ? reinterpret_cast<long long*>(static_cast<char*>(ws) + 1024) : nullptr;
  p.T = T; p.B = B; p.H = H; p.ND = ND; p.reverse_mask = reverse_mask; p.scale = 1.f; p.cap = cap;
  return persist_launch<true>(ta, tb0, tb1, p, smem, st);
}

}  // namespace ds2
