// Fused multi-tensor optimizer step: Adam + exponential moving average of the weights, one launch for all tensors.
// Reference: tf.train.AdamOptimizer(learning_rate).apply_gradients (src/deepSpeech_train.py:430, :457-458) and
// tf.train.ExponentialMovingAverage(moving_avg_decay, global_step).apply(trainables) (:461-462); the learning rate is
// tf.train.exponential_decay(..., staircase=True) (:231-250), computed on the host and passed in.
//   m = b1 m + (1-b1) g ;  v = b2 v + (1-b2) g^2 ;  w -= lr_t m / (sqrt(v) + eps),  lr_t = lr sqrt(1-b2^t)/(1-b1^t)
//   shadow -= (1 - decay) (shadow - w)            (with the UPDATED w; decay = min(d, (1+n)/(10+n)) from the host)
// g is multiplied by grad_scale first (1/world size when the all-reduce summed).  HBM-bound: 20 B read + 16 B
// written per parameter.
#include "common.cuh"

namespace ds2 {

static constexpr int OPT_CHUNK = 4096;          // elements per block iteration
static constexpr int OPT_COLS = 7;              // table row: w, g, m, v, shadow pointers, numel, first chunk

__global__ void __launch_bounds__(256)
adam_ema_kernel(const long long* __restrict__ table, int n_tensors, int n_chunks, float lr_t, float b1, float b2, float eps,
                float ema_keep, float grad_scale) {
  __shared__ int s_t;
  for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    if (threadIdx.x == 0) {
      int lo = 0, hi = n_tensors - 1;           // last tensor whose first chunk <= c
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (table[mid * OPT_COLS + 6] <= c) lo = mid; else hi = mid - 1;
      }
      s_t = lo;
    }
    __syncthreads();
    const long long* row = table + s_t * OPT_COLS;
    float* w = reinterpret_cast<float*>(row[0]);
    const float* g = reinterpret_cast<const float*>(row[1]);
    float* m = reinterpret_cast<float*>(row[2]);
    float* v = reinterpret_cast<float*>(row[3]);
    float* sh = reinterpret_cast<float*>(row[4]);
    const long long n = row[5];
    const long long i0 = (long long)(c - row[6]) * OPT_CHUNK;
    const long long i1 = i0 + OPT_CHUNK < n ? i0 + OPT_CHUNK : n;
    for (long long i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
      const float gi = g[i] * grad_scale;
      const float mi = b1 * m[i] + (1.f - b1) * gi;
      const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
      const float wi = w[i] - lr_t * mi / (sqrtf(vi) + eps);
      m[i] = mi;
      v[i] = vi;
      w[i] = wi;
      if (sh) sh[i] = sh[i] - (1.f - ema_keep) * (sh[i] - wi);
    }
    __syncthreads();
  }
}

}  // namespace ds2

using namespace ds2;

extern "C" int ds2_adam_ema_chunk(void) { return OPT_CHUNK; }

extern "C" int ds2_adam_ema_step(const void* table, int n_tensors, int n_chunks, float lr, float beta1, float beta2, float eps,
                                 long step, float ema_decay, float grad_scale, void* stream) {
  DS2_REQUIRE(table && n_tensors > 0 && n_chunks > 0 && step >= 1, DS2_ERR_BAD_ARG, "ds2_adam_ema_step: bad argument");
  const double t = (double)step;
  const float lr_t = (float)((double)lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t)));
  int grid = n_chunks < num_sms() * 8 ? n_chunks : num_sms() * 8;
  adam_ema_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const long long*>(table), n_tensors, n_chunks,
                                                                       lr_t, beta1, beta2, eps, ema_decay, grad_scale);
  DS2_LAUNCH_CHECK("adam_ema_kernel");
  count_launch();
  return DS2_OK;
}
