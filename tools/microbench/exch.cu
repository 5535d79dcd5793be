// Exchange-protocol micro-benchmark for the persistent recurrence (round 2).
// Reproduces ONLY the inter-CTA part of a recurrence timestep with the real launch geometry - grid (4 K-ranks, 14 unit
// tiles, 2 directions), clusters of 4, one "instance" of 16 utterances per CTA - and no tensor-core work:
//   each step a CTA (tile mt, rank r) PUBLISHES its 32 units x 16 utterances (1 KB, 128 threads x 8 B), then WAITS until
//   the 14 producers of its K slice have published, then FETCHES the 14 KB slice image into shared memory.
// Variants (argv[1] = mode):
//   0  round-1 protocol: rows in the [t][b][H] tensor, bar.sync + one red on a per-direction counter, one poller,
//      bar.sync, 7 x ld.relaxed.gpu.v4 per thread, st.shared
//   1  as 0 with ld.global.cg
//   2  image ring + per-direction counter (red by one thread after bar.sync) + ONE cp.async.bulk of 14 KB issued by the poller
//   3  image ring + PUSH counters: every warp of a producer adds 1 to the counter line of each of the 14 consumers of its
//      K slice (no CTA barrier); consumer polls its own line, then one bulk copy
//   4  as 3 but 7 x ld.relaxed.gpu.v4 per thread from the image instead of the bulk copy (after a named barrier)
//   5  as 3, no counters at all: the poller re-issues the bulk copy until no sentinel is left (pure data polling)
// argv[2] = emulated compute per step in cycles (spin), argv[3] = steps.  Every loaded value is checked (value = step tag).
#include <cstdio>
#include <cstdlib>

#include "../../deepspeech_b200/csrc/common.cuh"

using namespace ds2;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

static constexpr int KS = 4, TILES = 14, ND = 2, H = 1760, KSL = 448, BP = 16, RING = 8;
static constexpr int IMG_BYTES = KSL * BP * 2;   // 14336

__device__ __forceinline__ uint32_t ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void st_relaxed_b64(void* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint4 ldx(const void* p, int cg) {
  uint4 v;
  if (cg) asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  else asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void bulk_g2s(uint32_t sdst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sdst), "l"(gsrc), "r"(bytes), "r"(bar) : "memory");
}

struct P {
  unsigned short* rows;        // [T][ND][BP][H]  (modes 0,1)
  unsigned char* ring;         // [ND][RING][KS][IMG_BYTES]
  unsigned int* ctr;           // per-direction counters (32 uints apart) / per-consumer lines
  long long* out;              // per-CTA cycles
  unsigned int* bad;           // mismatches
  long long* stamps;           // optional stage stamps of CTA 0
  int T, mode, delay;
};

__global__ void __cluster_dims__(KS, 1, 1) __launch_bounds__(160, 1) exch_kernel(P p) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar[2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ctarank();
  const int mt = blockIdx.y, d = blockIdx.z;
  const int mode = p.mode;
  const bool image = mode >= 2;
  if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 128); fence_barrier_init(); }
  __syncthreads();
  // rows this CTA finishes: units [mt*128 + rank*32, +32) -> K slice ks_p = that / 448, offset inside it
  const int u0 = mt * 128 + (int)rank * 32;
  const bool produces = u0 < H;
  const int ks_p = u0 / KSL, uoff = u0 - ks_p * KSL;
  // consumers of my units: every tile's CTA with rank ks_p; consumer line index = (d*TILES + tile)*KS + rank
  unsigned int* dir_ctr = p.ctr + d * 32;
  unsigned int* my_line = p.ctr + 64 + ((d * TILES + mt) * KS + rank) * 32;
  const int nprod = (KSL * ((int)rank + 1) <= H ? KSL : H - KSL * (int)rank) / 32;   // producers of my K slice: 14,14,14,13
  long long t_start = 0;
  unsigned int nbad = 0;
  uint32_t ph = 0;
  const bool stamp = p.stamps && blockIdx.x + blockIdx.y + blockIdx.z == 0;
  for (int step = 0; step < p.T; ++step) {
    if (step == 2 && threadIdx.x == 0) t_start = clock64();
    if (stamp && threadIdx.x == 0) p.stamps[step * 8 + 0] = clock64();
    const int slot = step % RING;
    if (warp >= 1) {
      // ---------------- publish (epilogue role): thread = (utterance bl, 4-unit group jg)
      const int et = threadIdx.x - 32, bl = et % BP, jg = et / BP;
      const unsigned long long tag = (unsigned long long)(step + 1) * 0x0001000100010001ull;
      if (produces && mode != 6) {
        if (!image) {
          unsigned short* dst = p.rows + (((size_t)step * ND + d) * BP + bl) * H + u0 + 4 * jg;
          *reinterpret_cast<unsigned long long*>(dst) = tag;
        } else {
          unsigned char* img = p.ring + ((size_t)(d * RING + slot) * KS + ks_p) * IMG_BYTES;
          const int k = uoff + 4 * jg;
          const int off = (k >> 6) * (BP * 128) + bl * 128 + ((((k & 63) >> 3) ^ (bl & 7)) << 4) + (k & 7) * 2;
          st_relaxed_b64(img + off, tag);
          // rolling refill: the slot two steps ahead goes back to the sentinel
          unsigned char* img2 = p.ring + ((size_t)(d * RING + (step + 2) % RING) * KS + ks_p) * IMG_BYTES;
          st_relaxed_b64(img2 + off, ~0ull);
        }
      }
      if (mode <= 2) {
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0 && produces) red_relaxed_gpu_add(dir_ctr, 1u);
      } else if (mode <= 4) {
        __syncwarp();
        if (produces && lane < TILES) red_relaxed_gpu_add(p.ctr + 64 + ((d * TILES + lane) * KS + ks_p) * 32, 1u);
      }
      if (stamp && threadIdx.x == 32) p.stamps[step * 8 + 1] = clock64();
    }
    // ---------------- acquire + fetch
    const unsigned int need_dir = 55u * (unsigned)(step + 1);          // 55 producing CTAs per direction (H = 1760)
    const unsigned int need_mine = 4u * (unsigned)nprod * (unsigned)(step + 1);
    if (mode <= 1) {
      if (warp >= 1) {
        const int et = threadIdx.x - 32;
        if (et == 0) { while (ld_relaxed_gpu_u32(dir_ctr) < need_dir) {} }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const unsigned short* rows = p.rows + ((size_t)step * ND + d) * BP * H + (size_t)rank * KSL;
        uint4 v[7];
#pragma unroll
        for (int m = 0; m < 7; ++m) {
          const int u = et + 128 * m, ub = u / 56, uc = u - ub * 56;
          v[m] = make_uint4(0, 0, 0, 0);
          if ((int)rank * KSL + uc * 8 < H) v[m] = ldx(rows + ub * H + uc * 8, mode);
        }
#pragma unroll
        for (int m = 0; m < 7; ++m) {
          const int u = et + 128 * m, ub = u / 56, uc = u - ub * 56;
          if ((int)rank * KSL + uc * 8 < H) {
            while (unit_stale(v[m])) v[m] = ldx(rows + ub * H + uc * 8, mode);
            if ((v[m].x & 0xFFFF) != (unsigned)(step + 1)) ++nbad;
          }
        }
#pragma unroll
        for (int m = 0; m < 7; ++m) {
          const int u = et + 128 * m, ub = u / 56, uc = u - ub * 56;
          sts_v4(smem_u32(smem) + (uc >> 3) * (BP * 128) + ub * 128 + (((uc & 7) ^ (ub & 7)) << 4), v[m]);
        }
        fence_proxy_async();
        mbar_arrive(&bar[1]);
      } else {
        mbar_wait(&bar[1], ph, 1);
      }
    } else if (mode == 4) {
      if (warp >= 1) {
        const int et = threadIdx.x - 32;
        if (lane == 0) { while (ld_relaxed_gpu_u32(my_line) < need_mine) {} }
        __syncwarp();
        const unsigned char* img = p.ring + ((size_t)(d * RING + slot) * KS + rank) * IMG_BYTES;
        uint4 v[7];
#pragma unroll
        for (int m = 0; m < 7; ++m) {
          const int off = (et + 128 * m) * 16;
          v[m] = ldx(img + off, 0);
        }
#pragma unroll
        for (int m = 0; m < 7; ++m) {
          const int off = (et + 128 * m) * 16;
          const bool live = (int)rank * KSL + ((off >> 11) * 64) + ((((off >> 4) & 7) ^ ((off >> 7) & 7)) << 3) < H;
          while (live && unit_stale(v[m])) v[m] = ldx(img + off, 0);
          if (live && (v[m].x & 0xFFFF) != (unsigned)(step + 1)) ++nbad;
          sts_v4(smem_u32(smem) + off, v[m]);
        }
        fence_proxy_async();
        mbar_arrive(&bar[1]);
      } else {
        mbar_wait(&bar[1], ph, 1);
      }
    } else {
      // bulk copy by the control warp
      if (warp == 0) {
        const unsigned char* img = p.ring + ((size_t)(d * RING + slot) * KS + rank) * IMG_BYTES;
        if (lane == 0) {
          if (mode == 2) { while (ld_relaxed_gpu_u32(dir_ctr) < need_dir) {} }
          else if (mode == 3) { while (ld_relaxed_gpu_u32(my_line) < need_mine) {} }
        }
        __syncwarp();
        if (stamp && lane == 0) p.stamps[step * 8 + 2] = clock64();
        bool again = true;
        int ncopy = 0;
        while (again) {
          ++ncopy;
          if (lane == 0) {
            mbar_arrive_expect_tx(&bar[0], IMG_BYTES);
            bulk_g2s(smem_u32(smem), img, IMG_BYTES, smem_u32(&bar[0]));
          }
          mbar_wait(&bar[0], ph, 2);
          ph ^= 1;
          if (stamp && lane == 0) p.stamps[step * 8 + 3] = clock64();
          // validate: 32 lanes x 28 units
          bool stale = false;
          unsigned int wrong = 0;
#pragma unroll 4
          for (int m = 0; m < 28; ++m) {
            const int off = (lane + 32 * m) * 16;
            uint4 v = *reinterpret_cast<const uint4*>(smem + off);
            const bool live = (int)rank * KSL + ((off >> 11) * 64) + ((((off >> 4) & 7) ^ ((off >> 7) & 7)) << 3) < H;
            if (live) {
              stale |= unit_stale(v);
              wrong += (v.x & 0xFFFF) != (unsigned)(step + 1);
            }
          }
          again = mode != 6 && __any_sync(0xffffffffu, stale);
          if (!again && mode != 6) nbad += wrong;
        }
        if (stamp && lane == 0) { p.stamps[step * 8 + 4] = clock64(); p.stamps[step * 8 + 5] = ncopy; }
      }
    }
    if (mode <= 1 || mode == 4) ph ^= 1;
    __syncthreads();
    if (p.delay > 0) {
      const long long t0 = clock64();
      while (clock64() - t0 < p.delay) {}
    }
  }
  if (threadIdx.x == 0) p.out[(blockIdx.z * TILES + blockIdx.y) * KS + blockIdx.x] = clock64() - t_start;
  if (nbad) atomicAdd(p.bad, nbad);
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

int main(int argc, char** argv) {
  const int only = argc > 1 ? atoi(argv[1]) : -1;
  const int delay = argc > 2 ? atoi(argv[2]) : 0;
  const int T = argc > 3 ? atoi(argv[3]) : 400;
  P p = {};
  p.T = T; p.delay = delay;
  const size_t rows_bytes = (size_t)T * ND * BP * H * 2;
  const size_t ring_bytes = (size_t)ND * RING * KS * IMG_BYTES;
  CK(cudaMalloc(&p.rows, rows_bytes));
  CK(cudaMalloc(&p.ring, ring_bytes));
  CK(cudaMalloc(&p.ctr, 64 * 1024));
  CK(cudaMallocManaged(&p.out, 1024 * sizeof(long long)));
  CK(cudaMallocManaged(&p.bad, 4));
  CK(cudaMallocManaged(&p.stamps, (size_t)T * 8 * sizeof(long long)));
  CK(cudaFuncSetAttribute(exch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  const char* names[7] = {"round-1 protocol (counter, bar.sync, ld.relaxed.gpu)", "round-1 protocol with ld.global.cg",
                          "image ring + direction counter + one 14 KB bulk copy", "image ring + push counters (per-warp red to 14 consumers) + bulk copy",
                          "image ring + push counters + ld.relaxed.gpu per thread", "image ring, no counters: bulk copy re-issued until clean",
                          "bulk copy of the 14 KB image only (no publication, no wait)"};
  for (int mode = 0; mode < 7; ++mode) {
    if (only >= 0 && mode != only) continue;
    p.mode = mode;
    for (int rep = 0; rep < 2; ++rep) {
      CK(cudaMemset(p.rows, 0xFF, rows_bytes));
      CK(cudaMemset(p.ring, 0xFF, ring_bytes));
      CK(cudaMemset(p.ctr, 0, 64 * 1024));
      *p.bad = 0;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(KS, TILES, ND);
      cfg.blockDim = dim3(160);
      cfg.dynamicSmemBytes = 32 * 1024;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeCooperative;
      at[0].val.cooperative = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      CK(cudaLaunchKernelEx(&cfg, exch_kernel, p));
      CK(cudaDeviceSynchronize());
    }
    double mx = 0, mn = 1e30;
    for (int i = 0; i < KS * TILES * ND; ++i) { mx = p.out[i] > mx ? p.out[i] : mx; mn = p.out[i] < mn ? p.out[i] : mn; }
    printf("mode %d  %-76s: %6.0f cycles/step (delay %d -> exchange %6.0f), mismatches %u\n", mode, names[mode], mx / (T - 2), delay,
           mx / (T - 2) - delay, *p.bad);
    if (mode == 2 || mode == 3 || mode == 5 || mode == 6) {
      double g[5] = {0, 0, 0, 0, 0}, nc = 0;
      for (int t = 10; t < T - 1; ++t) {
        const long long* s = p.stamps + t * 8;
        g[0] += s[1] - s[0]; g[1] += s[2] - s[0]; g[2] += s[3] - s[2]; g[3] += s[4] - s[3]; g[4] += p.stamps[(t + 1) * 8] - s[4]; nc += s[5];
      }
      const double n = T - 11;
      printf("        CTA0: publish issued %.0f | step start -> poll done %.0f | bulk copy (last) %.0f | validate %.0f | to next step %.0f | copies/step %.2f\n",
             g[0] / n, g[1] / n, g[2] / n, g[3] / n, g[4] / n, nc / n);
    }
  }
  return 0;
}
