// Shape-generic fp32 recurrences (one small kernel per time step, all directions per launch).
// Exact-arithmetic path for shapes the persistent tcgen05 kernel does not take, and its parity
// baseline.  Cell math: src/custom_ops.py:48-72 (CustomRNNCell2); masking / reversal semantics of
// tf.nn.bidirectional_dynamic_rnn as used at src/custom_ops.py:91-93.
#include "common.cuh"

namespace ds2 {

// ------------------------------------------------------------------ clipped-ReLU RNN
__global__ void __launch_bounds__(256)
rnn_relu_step_fwd_kernel(const float* __restrict__ wx, float scale, const float* __restrict__ u, const float* __restrict__ bias,
                         const int* __restrict__ lens, float* __restrict__ hs, int T, int B, int H, int ND, int step,
                         int reverse_mask, float cap) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int d = blockIdx.z, b = blockIdx.y, j = blockIdx.x * 8 + warp;
  if (j >= H) return;
  const int rev = (reverse_mask >> d) & 1;
  const int t = rev ? T - 1 - step : step;
  const int tp = rev ? t + 1 : t - 1;
  float acc = 0.f;
  if (step > 0) {
    const float* hp = hs + (((size_t)d * T + tp) * B + b) * H;
    const float* ur = u + ((size_t)d * H + j) * H;
    for (int k = lane; k < H; k += 32) acc = fmaf(ur[k], hp[k], acc);
  }
  acc = warp_sum(acc);
  if (lane == 0) {
    const float z = scale * wx[((size_t)t * B + b) * ND * H + (size_t)d * H + j] + acc + bias[d * H + j];
    const float h = (t < lens[b]) ? fminf(fmaxf(z, 0.f), cap) : 0.f;
    hs[(((size_t)d * T + t) * B + b) * H + j] = h;
  }
}

__global__ void __launch_bounds__(256)
rnn_relu_step_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ u, const int* __restrict__ lens,
                         const float* __restrict__ hs, float* __restrict__ dz, int T, int B, int H, int ND, int step,
                         int reverse_mask, float cap) {
  const int d = blockIdx.z, b = blockIdx.y, k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= H) return;
  const int rev = (reverse_mask >> d) & 1;
  const int t = rev ? step : T - 1 - step;
  const int tn = rev ? t - 1 : t + 1;
  float dh = dout[((size_t)t * B + b) * H + k];
  if (step > 0) {
    const float* dzn = dz + ((size_t)tn * B + b) * ND * H + (size_t)d * H;
    const float* uc = u + (size_t)d * H * H + k;
    float acc = 0.f;
    for (int j = 0; j < H; ++j) acc = fmaf(uc[(size_t)j * H], dzn[j], acc);
    dh += acc;
  }
  const float h = hs[(((size_t)d * T + t) * B + b) * H + k];
  dz[((size_t)t * B + b) * ND * H + (size_t)d * H + k] = (t < lens[b] && h > 0.f && h < cap) ? dh : 0.f;
}

// ------------------------------------------------------------------ GRU (gate order r, z, n)
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

// gates [ND][T][B][4H] = r | z | n | (U_n h + bu_n)
__global__ void __launch_bounds__(256)
gru_step_fwd_kernel(const float* __restrict__ wx, float scale, const float* __restrict__ u, const float* __restrict__ bw,
                    const float* __restrict__ bu, const int* __restrict__ lens, float* __restrict__ hs, float* __restrict__ gates,
                    int T, int B, int H, int ND, int step, int reverse_mask) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int d = blockIdx.z, b = blockIdx.y, j = blockIdx.x * 8 + warp;
  if (j >= H) return;
  const int rev = (reverse_mask >> d) & 1;
  const int t = rev ? T - 1 - step : step;
  const int tp = rev ? t + 1 : t - 1;
  float ar = 0.f, az = 0.f, an = 0.f, hprev = 0.f;
  if (step > 0) {
    const float* hp = hs + (((size_t)d * T + tp) * B + b) * H;
    const float* ur = u + ((size_t)d * 3 * H + j) * H;
    for (int k = lane; k < H; k += 32) {
      const float hv = hp[k];
      ar = fmaf(ur[k], hv, ar);
      az = fmaf(ur[(size_t)H * H + k], hv, az);
      an = fmaf(ur[(size_t)2 * H * H + k], hv, an);
    }
    hprev = hp[j];
  }
  ar = warp_sum(ar); az = warp_sum(az); an = warp_sum(an);
  if (lane == 0) {
    const float* wxr = wx + ((size_t)t * B + b) * ND * 3 * H + (size_t)d * 3 * H;
    const float* bwd = bw + (size_t)d * 3 * H;
    const float* bud = bu + (size_t)d * 3 * H;
    const float r = sigmoidf_(scale * wxr[j] + bwd[j] + ar + bud[j]);
    const float z = sigmoidf_(scale * wxr[H + j] + bwd[H + j] + az + bud[H + j]);
    const float ghn = an + bud[2 * H + j];
    const float n = tanhf(scale * wxr[2 * H + j] + bwd[2 * H + j] + r * ghn);
    const bool valid = t < lens[b];
    const float h = valid ? (1.f - z) * n + z * hprev : 0.f;
    hs[(((size_t)d * T + t) * B + b) * H + j] = h;
    float* g = gates + (((size_t)d * T + t) * B + b) * 4 * H;
    g[j] = r; g[H + j] = z; g[2 * H + j] = n; g[3 * H + j] = ghn;
  }
}

// dgi/dgh [T][B][ND*3H]; dhz [ND][B][H] carries dh*z to the next (earlier) step.
__global__ void __launch_bounds__(256)
gru_step_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ u, const int* __restrict__ lens,
                    const float* __restrict__ hs, const float* __restrict__ gates, float* __restrict__ dgi,
                    float* __restrict__ dgh, float* __restrict__ dhz, int T, int B, int H, int ND, int step, int reverse_mask) {
  const int d = blockIdx.z, b = blockIdx.y, k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= H) return;
  const int rev = (reverse_mask >> d) & 1;
  const int t = rev ? step : T - 1 - step;
  const int tn = rev ? t - 1 : t + 1;      // processed in the previous backward step
  const int tp = rev ? t + 1 : t - 1;      // previous in forward order
  float dh = dout[((size_t)t * B + b) * H + k];
  float* carry = dhz + ((size_t)d * B + b) * H + k;
  if (step > 0) {
    const float* dn_ = dgh + ((size_t)tn * B + b) * ND * 3 * H + (size_t)d * 3 * H;
    const float* uc = u + (size_t)d * 3 * H * H + k;
    float acc = *carry;
    for (int j = 0; j < 3 * H; ++j) acc = fmaf(uc[(size_t)j * H], dn_[j], acc);
    dh += acc;
  }
  const bool valid = t < lens[b];
  if (!valid) dh = 0.f;
  const float* g = gates + (((size_t)d * T + t) * B + b) * 4 * H;
  const float r = g[k], z = g[H + k], n = g[2 * H + k], ghn = g[3 * H + k];
  const bool has_prev = rev ? (tp < T) : (tp >= 0);
  const float hprev = has_prev ? hs[(((size_t)d * T + tp) * B + b) * H + k] : 0.f;
  const float dn = dh * (1.f - z) * (1.f - n * n);
  const float dzg = dh * (hprev - n) * z * (1.f - z);
  const float dr = dn * ghn * r * (1.f - r);
  float* gi = dgi + ((size_t)t * B + b) * ND * 3 * H + (size_t)d * 3 * H;
  float* gh = dgh + ((size_t)t * B + b) * ND * 3 * H + (size_t)d * 3 * H;
  gi[k] = dr; gi[H + k] = dzg; gi[2 * H + k] = dn;
  gh[k] = dr; gh[H + k] = dzg; gh[2 * H + k] = dn * r;
  *carry = dh * z;
}

}  // namespace ds2

using namespace ds2;

namespace ds2 {
int rnn_relu_fwd_generic(const float* wx, float scale, const float* u, const float* bias, const int* lens, float* hs, int T,
                         int B, int H, int ND, int reverse_mask, float cap, cudaStream_t st) {
  dim3 grid((H + 7) / 8, B, ND);
  for (int step = 0; step < T; ++step) {
    rnn_relu_step_fwd_kernel<<<grid, 256, 0, st>>>(wx, scale, u, bias, lens, hs, T, B, H, ND, step, reverse_mask, cap);
  }
  DS2_LAUNCH_CHECK("rnn_relu_step_fwd_kernel");
  count_launch(T);
  return DS2_OK;
}
int rnn_relu_bwd_generic(const float* dout, const float* u, const int* lens, const float* hs, float* dz, int T, int B, int H,
                         int ND, int reverse_mask, float cap, cudaStream_t st) {
  dim3 grid((H + 127) / 128, B, ND);
  for (int step = 0; step < T; ++step) {
    rnn_relu_step_bwd_kernel<<<grid, 128, 0, st>>>(dout, u, lens, hs, dz, T, B, H, ND, step, reverse_mask, cap);
  }
  DS2_LAUNCH_CHECK("rnn_relu_step_bwd_kernel");
  count_launch(T);
  return DS2_OK;
}
}  // namespace ds2

extern "C" int ds2_gru_fwd(const float* wx, float scale, const void* u, const float* bw, const float* bu, const int* lens,
                           void* hs, void* gates, int dtype, int T, int B, int H, int ND, int reverse_mask, void* stream) {
  DS2_REQUIRE(dtype == DS2_F32, DS2_ERR_UNSUPPORTED, "ds2_gru_fwd: only the fp32 path is built (dtype=%d)", dtype);
  DS2_REQUIRE(wx && u && bw && bu && lens && hs && gates, DS2_ERR_BAD_ARG, "ds2_gru_fwd: null pointer");
  DS2_REQUIRE(T > 0 && B > 0 && H > 0 && (ND == 1 || ND == 2), DS2_ERR_BAD_SHAPE, "ds2_gru_fwd: bad shape");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid((H + 7) / 8, B, ND);
  for (int step = 0; step < T; ++step)
    gru_step_fwd_kernel<<<grid, 256, 0, st>>>(wx, scale, static_cast<const float*>(u), bw, bu, lens, static_cast<float*>(hs),
                                              static_cast<float*>(gates), T, B, H, ND, step, reverse_mask);
  DS2_LAUNCH_CHECK("gru_step_fwd_kernel");
  count_launch(T);
  return DS2_OK;
}

extern "C" int ds2_gru_bwd(const float* dout, const void* u, const int* lens, const void* hs, const void* gates, void* dgi,
                           void* dgh, float* dhz_scratch, int dtype, int T, int B, int H, int ND, int reverse_mask, void* stream) {
  DS2_REQUIRE(dtype == DS2_F32, DS2_ERR_UNSUPPORTED, "ds2_gru_bwd: only the fp32 path is built (dtype=%d)", dtype);
  DS2_REQUIRE(dout && u && lens && hs && gates && dgi && dgh && dhz_scratch, DS2_ERR_BAD_ARG, "ds2_gru_bwd: null pointer");
  DS2_REQUIRE(T > 0 && B > 0 && H > 0 && (ND == 1 || ND == 2), DS2_ERR_BAD_SHAPE, "ds2_gru_bwd: bad shape");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid((H + 127) / 128, B, ND);
  for (int step = 0; step < T; ++step)
    gru_step_bwd_kernel<<<grid, 128, 0, st>>>(dout, static_cast<const float*>(u), lens, static_cast<const float*>(hs),
                                              static_cast<const float*>(gates), static_cast<float*>(dgi), static_cast<float*>(dgh),
                                              dhz_scratch, T, B, H, ND, step, reverse_mask);
  DS2_LAUNCH_CHECK("gru_step_bwd_kernel");
  count_launch(T);
  return DS2_OK;
}
