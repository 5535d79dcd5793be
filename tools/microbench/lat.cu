// Latency micro-benchmarks behind the design of the persistent recurrence (round 2): how long the hand-offs a
// timestep is made of really take on this B200, in SM cycles at the clock the kernel runs at.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o lat lat.cu && ./lat
// Prints one line per experiment: cycles per one-way hand-off (ping-pong round trip / 2).
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../deepspeech_b200/csrc/common.cuh"

using namespace ds2;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void st_relaxed_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_b64(void* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_b64(const void* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_v4u(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int smid() {
  unsigned int r;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(r));
  return r;
}

// ---------------------------------------------------------------- (1) pointer chase through L2 (ld.relaxed.gpu / ld.cg)
__global__ void chase_kernel(const unsigned int* next, int n, int mode, long long* out) {
  unsigned int i = 0;
  long long t0 = clock64();
  for (int k = 0; k < n; ++k) {
    if (mode == 0) i = ld_relaxed_gpu_u32(next + i);
    else asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(i) : "l"(next + i) : "memory");
  }
  long long t1 = clock64();
  out[0] = t1 - t0;
  out[1] = i;
}

// ---------------------------------------------------------------- (2) flag ping-pong between two SMs
// mode 0: st.relaxed flag / poll        mode 1: red.relaxed.add flag / poll
// mode 2: "counter protocol": weak 8-byte data store, __syncwarp, red on a counter; peer polls the counter, then loads
//         the data (ld.relaxed v4) - what the round-1 kernel does per step (one thread instead of 128)
// mode 3: "LL": data carries its own sequence number (st.relaxed.b64), peer polls the data itself
__global__ void pingpong_kernel(unsigned int* flags, unsigned long long* data, int n, int mode, long long* out) {
  const int me = blockIdx.x;          // 0 or 1
  unsigned int* mine = flags + me * 64;
  unsigned int* theirs = flags + (1 - me) * 64;
  unsigned long long* dmine = data + me * 64;
  unsigned long long* dtheirs = data + (1 - me) * 64;
  if (threadIdx.x != 0) return;
  long long t0 = clock64();
  for (int i = 1; i <= n; ++i) {
    if (me == 0) {
      // send
      if (mode == 0) st_relaxed_u32(mine, i);
      else if (mode == 1) red_relaxed_gpu_add(mine, 1u);
      else if (mode == 2) { dmine[i & 7] = (unsigned long long)i * 0x100000001ull; red_relaxed_gpu_add(mine, 1u); }
      else st_relaxed_b64(dmine, (unsigned long long)i * 0x100000001ull);
      // receive
      if (mode <= 1) { while (ld_relaxed_gpu_u32(theirs) < (unsigned)i) {} }
      else if (mode == 2) {
        while (ld_relaxed_gpu_u32(theirs) < (unsigned)i) {}
        while ((unsigned)ld_relaxed_b64(dtheirs + (i & 7)) != (unsigned)i) {}
      } else { while ((unsigned)ld_relaxed_b64(dtheirs) != (unsigned)i) {} }
    } else {
      if (mode <= 1) { while (ld_relaxed_gpu_u32(theirs) < (unsigned)i) {} }
      else if (mode == 2) {
        while (ld_relaxed_gpu_u32(theirs) < (unsigned)i) {}
        while ((unsigned)ld_relaxed_b64(dtheirs + (i & 7)) != (unsigned)i) {}
      } else { while ((unsigned)ld_relaxed_b64(dtheirs) != (unsigned)i) {} }
      if (mode == 0) st_relaxed_u32(mine, i);
      else if (mode == 1) red_relaxed_gpu_add(mine, 1u);
      else if (mode == 2) { dmine[i & 7] = (unsigned long long)i * 0x100000001ull; red_relaxed_gpu_add(mine, 1u); }
      else st_relaxed_b64(dmine, (unsigned long long)i * 0x100000001ull);
    }
  }
  long long t1 = clock64();
  if (me == 0) { out[0] = t1 - t0; out[1] = smid(); } else out[2] = smid();
}

// ---------------------------------------------------------------- (3) all-to-all step barrier among G CTAs through L2
// every CTA: publish (mode 0: red on ONE shared counter, poll it; mode 1: LL - write own seq word, poll all G words,
// 32 lanes x G/32 words each) - the per-step pattern of a recurrence direction (G = 56)
__global__ void allgather_kernel(unsigned int* ctr, unsigned long long* words, int G, int n, int mode, long long* out) {
  const int me = blockIdx.x;
  const int lane = threadIdx.x;
  long long t0 = clock64();
  for (int i = 1; i <= n; ++i) {
    if (mode == 0) {
      if (lane == 0) {
        red_relaxed_gpu_add(ctr, 1u);
        while (ld_relaxed_gpu_u32(ctr) < (unsigned)(i * G)) {}
      }
      __syncwarp();
    } else {
      if (lane == 0) st_relaxed_b64(words + me * 16, (unsigned long long)i);
      for (int w = lane; w < G; w += 32) {
        while ((unsigned)ld_relaxed_b64(words + w * 16) < (unsigned)i) {}
      }
      __syncwarp();
    }
  }
  long long t1 = clock64();
  if (lane == 0) out[me] = t1 - t0;
}

// ---------------------------------------------------------------- (4) DSMEM st.async + mbarrier ping-pong inside a 2-CTA cluster
__device__ __forceinline__ uint32_t ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__global__ void __cluster_dims__(2, 1, 1) dsmem_kernel(int n, long long* out) {
  __shared__ __align__(16) uint32_t buf[4];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t rank = ctarank();
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (threadIdx.x == 0) {
    uint32_t rbuf, rbar;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rbuf) : "r"(smem_u32(buf)), "r"(1 - rank));
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rbar) : "r"(smem_u32(&bar)), "r"(1 - rank));
    long long t0 = clock64();
    uint32_t ph = 0;
    for (int i = 1; i <= n; ++i) {
      if (rank == 0) {
        mbar_arrive_expect_tx(&bar, 16);
        asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%2, %2, %2, %2}, [%1];" ::"r"(rbuf), "r"(rbar), "r"(i) : "memory");
        while (!mbar_try_wait(&bar, ph)) {}
      } else {
        mbar_arrive_expect_tx(&bar, 16);
        while (!mbar_try_wait(&bar, ph)) {}
        asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%2, %2, %2, %2}, [%1];" ::"r"(rbuf), "r"(rbar), "r"(i) : "memory");
      }
      ph ^= 1;
    }
    long long t1 = clock64();
    if (rank == 0) out[0] = t1 - t0;
  }
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// ---------------------------------------------------------------- (5) tcgen05.mma issue + completion latency, A in TMEM
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}
template <int N, int NMMA, bool A_SMEM>
__global__ void mma_kernel(int reps, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&slot, 512);
  // zero the operands (bf16 zeros: finite math)
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot;
  if (warp == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
    const uint32_t b0 = smem_u32(smem);
    const uint32_t a0 = smem_u32(smem) + 32768;
    uint32_t ph = 0;
    long long tot_issue = 0, tot_all = 0;
    for (int r = 0; r < reps; ++r) {
      long long t0 = clock64(), t1 = 0;
      if (elect_one_sync()) {
#pragma unroll
        for (int i = 0; i < NMMA; ++i) {
          const uint64_t bd = umma_smem_desc_sw128(b0 + (i >> 2) * (N * 128) + (i & 3) * 32, 16, 1024);
          if (A_SMEM) umma_bf16(tb + 448, umma_smem_desc_sw128(a0 + (i & 3) * 32, 16, 1024), bd, idesc, i > 0);
          else mma_ts(tb + 448, tb + (i * 8) % 384, bd, idesc, i > 0);
        }
        umma_commit(&bar);
        t1 = clock64();
      }
      __syncwarp();
      t1 = __shfl_sync(0xffffffffu, (unsigned long long)t1, __ffs(__ballot_sync(0xffffffffu, t1 != 0)) - 1);
      while (!mbar_try_wait(&bar, ph)) {}
      long long t2 = clock64();
      ph ^= 1;
      if (r > 0) { tot_issue += t1 - t0; tot_all += t2 - t0; }
    }
    if (threadIdx.x == 0) { out[0] = tot_issue / (reps - 1); out[1] = tot_all / (reps - 1); }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tb, 512); }
}

int main() {
  long long* out;
  CK(cudaMallocManaged(&out, 4096 * sizeof(long long)));
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  printf("device clock attr %d MHz (cycles below are SM clock64 cycles at whatever clock the GPU boosts to)\n", clk_khz / 1000);

  // (1) chase: 64 MB buffer (fits L2), stride permutation so that every load is a different line
  {
    const int n_el = 1 << 22;   // 16 MB of uint32: L2 resident
    std::vector<unsigned int> h(n_el);
    const unsigned int stride = 1031 * 32;   // lines apart, co-prime walk
    for (unsigned int i = 0; i < (unsigned)n_el; ++i) h[i] = (i + stride) % n_el;
    unsigned int* d;
    CK(cudaMalloc(&d, n_el * 4));
    CK(cudaMemcpy(d, h.data(), n_el * 4, cudaMemcpyHostToDevice));
    for (int mode = 0; mode < 2; ++mode) {
      chase_kernel<<<1, 1>>>(d, 2000, mode, out);   // warm (brings lines into L2)
      CK(cudaDeviceSynchronize());
      chase_kernel<<<1, 1>>>(d, 2000, mode, out);
      CK(cudaDeviceSynchronize());
      printf("L2 pointer chase (%s): %.0f cycles per dependent load\n", mode ? "ld.global.cg" : "ld.relaxed.gpu", (double)out[0] / 2000);
    }
    CK(cudaFree(d));
  }
  // (2) ping-pong
  {
    unsigned int* flags;
    unsigned long long* data;
    CK(cudaMalloc(&flags, 4096));
    CK(cudaMalloc(&data, 4096));
    const char* names[4] = {"st.relaxed flag -> poll", "red.add flag -> poll", "data st + red counter -> poll -> ld data (round-1 protocol)",
                            "LL: st.relaxed.b64 data+seq -> poll the data"};
    for (int mode = 0; mode < 4; ++mode) {
      for (int rep = 0; rep < 2; ++rep) {
        CK(cudaMemset(flags, 0, 4096));
        CK(cudaMemset(data, 0, 4096));
        pingpong_kernel<<<2, 32>>>(flags, data, 2000, mode, out);
        CK(cudaDeviceSynchronize());
      }
      printf("ping-pong SM%lld<->SM%lld %-62s: %.0f cycles one way\n", out[1], out[2], names[mode], (double)out[0] / 2000 / 2);
    }
    // (3) all-to-all
    for (int G : {4, 14, 56, 112}) {
      for (int mode = 0; mode < 2; ++mode) {
        CK(cudaMemset(flags, 0, 4096));
        unsigned long long* words;
        CK(cudaMalloc(&words, 112 * 128));
        CK(cudaMemset(words, 0, 112 * 128));
        allgather_kernel<<<G, 32>>>(flags, words, G, 1000, mode, out);
        CK(cudaDeviceSynchronize());
        double mx = 0;
        for (int i = 0; i < G; ++i) mx = out[i] > mx ? out[i] : mx;
        printf("step barrier among %3d CTAs (%s): %.0f cycles per step\n", G, mode ? "LL words, each CTA polls all" : "one counter, red + poll", mx / 1000);
        CK(cudaFree(words));
      }
    }
    CK(cudaFree(flags));
    CK(cudaFree(data));
  }
  // (4) DSMEM
  {
    dsmem_kernel<<<2, 32>>>(2000, out);
    CK(cudaDeviceSynchronize());
    printf("DSMEM st.async + mbarrier complete_tx ping-pong: %.0f cycles one way\n", (double)out[0] / 2000 / 2);
  }
  // (5) MMA
  {
#define RUN_MMA(N, NM, AS)                                                                                          \
  {                                                                                                                  \
    auto k = mma_kernel<N, NM, AS>;                                                                                  \
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));                            \
    k<<<1, 128, 100 * 1024>>>(50, out);                                                                              \
    CK(cudaDeviceSynchronize());                                                                                     \
    printf("tcgen05.mma M128 N%-3d K16 x %2d, A in %s: issue %lld cycles, issue -> commit observed %lld cycles\n", N, NM, \
           AS ? "smem" : "TMEM", out[0], out[1]);                                                                    \
  }
    RUN_MMA(16, 28, false);
    RUN_MMA(32, 28, false);
    RUN_MMA(16, 4, false);
    RUN_MMA(16, 1, false);
    RUN_MMA(16, 56, false);
    RUN_MMA(32, 56, false);
    RUN_MMA(16, 28, true);
    RUN_MMA(64, 28, false);
  }
  return 0;
}
