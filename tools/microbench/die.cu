// Bulk-copy latency by L2 home die (round 2, B200): one CTA copies 2 KB / 4 KB / 14 KB with cp.async.bulk from pages that
// are homed on its own die, on the other die, or mixed (pages classified first with dependent strong loads).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o die die.cu && ./die
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../deepspeech_b200/csrc/common.cuh"
using namespace ds2;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ int load_lat(const unsigned int* p, int reps) {
  unsigned int w = ld_relaxed_gpu_u32(p);
  int best = 1 << 30;
  for (int trial = 0; trial < 3; ++trial) {
    const long long t0 = clock64();
    for (int i = 0; i < reps; ++i) w += ld_relaxed_gpu_u32(p + (w >> 31));
    const long long t1 = clock64();
    best = min(best, (int)((t1 - t0) / reps) + (int)(w >> 31));
  }
  return best;
}
// pages: lat[i] = strong-load latency of 4 KB page i seen from this SM
__global__ void classify_pages(const unsigned int* buf, int npages, int* lat) {
  if (threadIdx.x == 0)
    for (int i = 0; i < npages; ++i) lat[i] = load_lat(buf + (size_t)i * 1024, 4);
}
// list[k] = page index of the k-th 4 KB piece to copy; bytes_per_piece <= 4096
__global__ void copy_lat(const unsigned char* buf, const int* list, int npieces, int bytes_per_piece, int reps, int* out) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
    long long best = 1ll << 60, sum = 0;
    uint32_t ph = 0;
    for (int r = 0; r < reps; ++r) {
      const long long t0 = clock64();
      mbar_arrive_expect_tx(&bar, npieces * bytes_per_piece);
      for (int k = 0; k < npieces; ++k)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sm + k * bytes_per_piece)), "l"(buf + (size_t)list[k] * 4096), "r"(bytes_per_piece), "r"(smem_u32(&bar)) : "memory");
      mbar_wait(&bar, ph, 1);
      ph ^= 1;
      const long long dt = clock64() - t0;
      if (r > 0) { sum += dt; best = dt < best ? dt : best; }
    }
    out[0] = (int)best;
    out[1] = (int)(sum / (reps - 1));
  }
}

int main() {
  const int npages = 256;
  unsigned char* buf;
  CK(cudaMalloc(&buf, (size_t)npages * 4096));
  CK(cudaMemset(buf, 0, (size_t)npages * 4096));
  int *lat, *list, *out;
  CK(cudaMallocManaged(&lat, npages * 4));
  CK(cudaMallocManaged(&list, 64 * 4));
  CK(cudaMallocManaged(&out, 8));
  CK(cudaFuncSetAttribute(copy_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  // a single-CTA grid lands on the same SM every time on an idle GPU: the classification holds for the copies
  classify_pages<<<1, 32>>>(reinterpret_cast<unsigned int*>(buf), npages, lat);
  CK(cudaDeviceSynchronize());
  std::vector<int> nearp, farp;
  int lo = 1 << 30, hi = 0;
  for (int i = 0; i < npages; ++i) { lo = lat[i] < lo ? lat[i] : lo; hi = lat[i] > hi ? lat[i] : hi; }
  printf("strong-load latency of the pages: %d .. %d clk\n", lo, hi);
  for (int i = 0; i < npages; ++i) (lat[i] < (lo + hi) / 2 ? nearp : farp).push_back(i);
  if (nearp.size() < 8 || farp.size() < 8) { printf("pages not separable\n"); return 1; }
  printf("%zu near pages (e.g. %d clk), %zu far pages (e.g. %d clk)\n", nearp.size(), lat[nearp[0]], farp.size(), lat[farp[0]]);
  struct Case { const char* name; int npieces, bytes, nfar; };
  const Case cases[] = {{"2 KB", 1, 2048, 0}, {"2 KB", 1, 2048, 1}, {"4 KB", 1, 4096, 0}, {"4 KB", 1, 4096, 1},
                        {"14 KB in 4 pieces", 4, 3584, 0}, {"14 KB in 4 pieces", 4, 3584, 4}, {"14 KB in 4 pieces", 4, 3584, 2},
                        {"28 KB in 8 pieces", 8, 3584, 0}, {"28 KB in 8 pieces", 8, 3584, 8}};
  for (const Case& c : cases) {
    for (int k = 0; k < c.npieces; ++k) list[k] = k < c.nfar ? farp[k] : nearp[k];
    copy_lat<<<1, 32, 64 * 1024>>>(buf, list, c.npieces, c.bytes, 200, out);
    CK(cudaDeviceSynchronize());
    printf("bulk copy %-18s %d of %d pieces on the far die: best %4d clk, mean %4d clk\n", c.name, c.nfar, c.npieces, out[0], out[1]);
  }
  return 0;
}
