// L2 home die experiments (round 2, B200): which SMs are near which addresses, and what a far-die counter costs.
// Step 1: every SM, one at a time, times atomics on byte 0 of a buffer (near ~300 clk, far ~700 clk) -> SM classes;
//         SM of block 0 times atomics on every 256-byte step of the buffer -> home-die pattern of the addresses.
// Step 2: G CTAs that all sit on the die of byte 0 run a counter barrier (red + poll) with the counter homed on the
//         same / on the other die.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o topo2 topo2.cu && ./topo2
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../deepspeech_b200/csrc/common.cuh"
using namespace ds2;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ int atom_lat(unsigned int* p, int reps) {
  unsigned int w = atomicAdd(p, 0u);
  int best = 1 << 30;
  for (int trial = 0; trial < 3; ++trial) {            // a chain of dependent atomics; the best of three chains
    const long long t0 = clock64();
    for (int i = 0; i < reps; ++i) w += atomicAdd(p + (w & 1), 0u);
    const long long t1 = clock64();
    best = min(best, (int)((t1 - t0) / reps) + (int)(w & 0));
  }
  return best;
}
__device__ __forceinline__ void my_turn(int slot) {
  const long long t0 = clock64();
  while (clock64() - t0 < (long long)slot * 30000) {}
}
__global__ void classify_kernel(unsigned int* buf, int nsteps, int* sm_lat, int* smid_out, int* addr_lat) {
  if (threadIdx.x != 0) return;
  unsigned int smid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  smid_out[blockIdx.x] = smid;
  my_turn(blockIdx.x);                       // one SM at a time: no contention on the line
  sm_lat[blockIdx.x] = atom_lat(buf, 8);
  if (blockIdx.x == 0) {
    my_turn(160);
    for (int i = 0; i < nsteps; ++i) addr_lat[i] = atom_lat(buf + i * 64, 4);
  }
}
__global__ void barrier_kernel(unsigned int* buf, int G, unsigned int* ticket, unsigned int* ctr, int reps, long long* out) {
  if (threadIdx.x != 0) return;
  my_turn(blockIdx.x);
  if (atom_lat(buf, 8) >= 500) return;       // only SMs on the die of byte 0
  const unsigned int t = atomicAdd(ticket, 1u);
  if ((int)t >= G) return;
  red_relaxed_gpu_add(ctr + 32, 1u);
  while (ld_relaxed_gpu_u32(ctr + 32) < (unsigned)G) {}
  const long long t0 = clock64();
  for (int i = 1; i <= reps; ++i) {
    red_relaxed_gpu_add(ctr, 1u);
    while (ld_relaxed_gpu_u32(ctr) < (unsigned)(i * G)) {}
  }
  out[t] = clock64() - t0;
}

int main() {
  const int nb = 148, nsteps = 1024;
  unsigned int* buf;
  CK(cudaMalloc(&buf, 512 * 1024));
  CK(cudaMemset(buf, 0, 512 * 1024));
  int *sm_lat, *smid, *addr_lat;
  CK(cudaMallocManaged(&sm_lat, nb * 4));
  CK(cudaMallocManaged(&smid, nb * 4));
  CK(cudaMallocManaged(&addr_lat, nsteps * 4));
  CK(cudaFuncSetAttribute(classify_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CK(cudaFuncSetAttribute(barrier_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  classify_kernel<<<nb, 32, 200 * 1024>>>(buf, nsteps, sm_lat, smid, addr_lat);   // 200 KB of smem: one CTA per SM
  CK(cudaDeviceSynchronize());
  int n0 = 0, ref_lat = 0;
  printf("atomic round trip to byte 0, per SM (smid:clk):");
  for (int i = 0; i < nb; ++i) {
    n0 += sm_lat[i] < 500;
    if (i % 16 == 0) printf("\n ");
    printf(" %3d:%-4d", smid[i], sm_lat[i]);
  }
  ref_lat = sm_lat[0];
  printf("\nSMs near byte 0: %d of %d\n", n0, nb);
  const bool ref_near = ref_lat < 500;
  printf("home die per 256 B of the first 96 KB, seen from SM %d ('.' die of byte 0, '#' the other):\n", smid[0]);
  for (int i = 0; i < 384; ++i) {
    putchar(((addr_lat[i] < 500) == ref_near) ? '.' : '#');
    if (i % 96 == 95) putchar('\n');
  }
  int same_blk = -1, other_blk = -1;
  for (int b = 8; b < nsteps / 16 && (same_blk < 0 || other_blk < 0); ++b) {    // 4 KB blocks
    int same = 0;
    for (int j = 0; j < 16; ++j) same += ((addr_lat[b * 16 + j] < 500) == ref_near);
    if (same == 16 && same_blk < 0) same_blk = b;
    if (same == 0 && other_blk < 0) other_blk = b;
  }
  printf("4 KB block %d is homed with byte 0, block %d on the other die\n", same_blk, other_blk);
  unsigned int* ticket;
  long long* out;
  CK(cudaMallocManaged(&ticket, 4));
  CK(cudaMallocManaged(&out, 256 * 8));
  for (int G : {14, 56}) {
    for (int far = 0; far < 2; ++far) {
      unsigned int* ctr = buf + (size_t)(far ? other_blk : same_blk) * 1024;
      CK(cudaMemset(ctr, 0, 512));
      *ticket = 0;
      for (int i = 0; i < 256; ++i) out[i] = 0;
      barrier_kernel<<<nb, 32, 200 * 1024>>>(buf, G, ticket, ctr, 500, out);
      CK(cudaDeviceSynchronize());
      double mx = 0;
      for (int i = 0; i < G; ++i) mx = out[i] > mx ? out[i] : mx;
      printf("%2d CTAs on the die of byte 0 (%u took part), counter barrier, counter homed on the %s die: %.0f cycles per step\n", G,
             *ticket < (unsigned)G ? *ticket : (unsigned)G, far ? "OTHER" : "SAME ", mx / 500);
    }
  }
  return 0;
}
