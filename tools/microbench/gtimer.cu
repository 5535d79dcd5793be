// resolution of %globaltimer on B200: distinct values seen by a thread reading it back to back
#include <cstdio>
__global__ void k(unsigned long long* out, int n) {
  if (threadIdx.x) return;
  for (int i = 0; i < n; ++i) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    out[blockIdx.x * n + i] = t;
  }
}
int main() {
  const int n = 4096;
  unsigned long long* o;
  cudaMallocManaged(&o, 2 * n * 8);
  k<<<2, 32>>>(o, n);
  cudaDeviceSynchronize();
  for (int b = 0; b < 2; ++b) {
    int changes = 0; unsigned long long mind = ~0ull, maxd = 0;
    for (int i = 1; i < n; ++i) { unsigned long long d = o[b * n + i] - o[b * n + i - 1]; if (d) { ++changes; mind = d < mind ? d : mind; maxd = d > maxd ? d : maxd; } }
    printf("block %d: %d reads over %llu ns, %d value changes, step min %llu ns max %llu ns\n", b, n, o[b * n + n - 1] - o[b * n], changes, mind, maxd);
  }
  return 0;
}
