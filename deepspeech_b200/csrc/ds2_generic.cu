// Shape-generic fp32 kernels: strided GEMM, direct convolution (fwd / dgrad / wgrad),
// batch-stat BN + clipped ReLU (fwd / bwd), element-wise helpers.  These are the exact-arithmetic
// path for small or unaligned shapes and the parity baseline for the tcgen05 kernels.
#include "common.cuh"

namespace ds2 {

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

// ------------------------------------------------------------------ strided GEMM
// C[m,n] = alpha * sum_k A[m*sam + k*sak] * B[n*sbn + k*sbk] + beta * C[m,n] + bias[n]
template <typename TA, typename TB, typename TC>
__global__ void __launch_bounds__(256)
gemm_generic_kernel(const TA* __restrict__ A, long sam, long sak, const TB* __restrict__ B, long sbn, long sbk,
                    TC* __restrict__ C, long ldc, int M, int N, int K, float alpha, float beta,
                    const float* __restrict__ bias) {
  __shared__ float As[16][65];
  __shared__ float Bs[16][65];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      int r, kk;
      if (sak == 1) { kk = i % 16; r = i / 16; } else { r = i % 64; kk = i / 64; }
      const int m = m0 + r, k = k0 + kk;
      As[kk][r] = (m < M && k < K) ? ldf(A + (long)m * sam + (long)k * sak) : 0.f;
    }
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      int r, kk;
      if (sbk == 1) { kk = i % 16; r = i / 16; } else { r = i % 64; kk = i / 64; }
      const int n = n0 + r, k = k0 + kk;
      Bs[kk][r] = (n < N && k < K) ? ldf(B + (long)n * sbn + (long)k * sbk) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty + 16 * i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx + 16 * j;
      if (n >= N) continue;
      float v = alpha * acc[i][j];
      if (bias) v += bias[n];
      if (beta != 0.f) v += beta * ldf(C + (long)m * ldc + n);
      stf(C + (long)m * ldc + n, v);
    }
  }
}

template <typename TA, typename TB, typename TC>
static int gemm_generic_launch(const void* A, long sam, long sak, const void* B, long sbn, long sbk, void* C, long ldc, int M,
                               int N, int K, float alpha, float beta, const float* bias, cudaStream_t st) {
  dim3 grid((N + 63) / 64, (M + 63) / 64);
  gemm_generic_kernel<TA, TB, TC><<<grid, 256, 0, st>>>(static_cast<const TA*>(A), sam, sak, static_cast<const TB*>(B), sbn, sbk,
                                                        static_cast<TC*>(C), ldc, M, N, K, alpha, beta, bias);
  DS2_LAUNCH_CHECK("gemm_generic_kernel");
  count_launch();
  return DS2_OK;
}

// ------------------------------------------------------------------ direct convolution, channel-last
// x [B,Ti,Fi,Ci]  w [KT,KF,Ci,Co] (HWIO, as tf.nn.conv2d)  y [B,To,Fo,Co]; VALID, strides (st,sf).
struct ConvShape {
  int B, Ti, Fi, Ci, To, Fo, Co, KT, KF, st, sf;
};

template <typename TX>
__global__ void __launch_bounds__(256)
conv_fwd_kernel(const TX* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ y,
                ConvShape s) {
  const long total = (long)s.B * s.To * s.Fo * s.Co;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int co = i % s.Co;
    long r = i / s.Co;
    const int fo = r % s.Fo; r /= s.Fo;
    const int to = r % s.To;
    const int b = r / s.To;
    float acc = bias ? bias[co] : 0.f;
    for (int kt = 0; kt < s.KT; ++kt) {
      const TX* xr = x + (((long)b * s.Ti + (to * s.st + kt)) * s.Fi + fo * s.sf) * s.Ci;
      const float* wr = w + (long)kt * s.KF * s.Ci * s.Co + co;
      for (int q = 0; q < s.KF * s.Ci; ++q) acc = fmaf(ldf(xr + q), wr[(long)q * s.Co], acc);
    }
    y[i] = acc;
  }
}

__global__ void __launch_bounds__(256)
conv_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w, float* __restrict__ dx, ConvShape s) {
  const long total = (long)s.B * s.Ti * s.Fi * s.Ci;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int ci = i % s.Ci;
    long r = i / s.Ci;
    const int fi = r % s.Fi; r /= s.Fi;
    const int ti = r % s.Ti;
    const int b = r / s.Ti;
    float acc = 0.f;
    for (int kt = 0; kt < s.KT; ++kt) {
      const int tn = ti - kt;
      if (tn < 0 || tn % s.st) continue;
      const int to = tn / s.st;
      if (to >= s.To) continue;
      for (int kf = 0; kf < s.KF; ++kf) {
        const int fn = fi - kf;
        if (fn < 0 || fn % s.sf) continue;
        const int fo = fn / s.sf;
        if (fo >= s.Fo) continue;
        const float* dyr = dy + (((long)b * s.To + to) * s.Fo + fo) * s.Co;
        const float* wr = w + (((long)kt * s.KF + kf) * s.Ci + ci) * s.Co;
        for (int co = 0; co < s.Co; ++co) acc = fmaf(dyr[co], wr[co], acc);
      }
    }
    dx[i] = acc;
  }
}

// dw must be zeroed by the caller (host wrapper does it); blockIdx.y splits the (b,to) rows.
template <typename TX>
__global__ void __launch_bounds__(256)
conv_wgrad_kernel(const TX* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dw, float* __restrict__ dbias,
                  ConvShape s, int rows_per_block) {
  const int nw = s.KT * s.KF * s.Ci * s.Co;
  const int row0 = blockIdx.y * rows_per_block;
  const int row1 = min(row0 + rows_per_block, s.B * s.To);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nw + s.Co; i += gridDim.x * blockDim.x) {
    float acc = 0.f;
    if (i < nw) {
      const int co = i % s.Co;
      int r = i / s.Co;
      const int ci = r % s.Ci; r /= s.Ci;
      const int kf = r % s.KF;
      const int kt = r / s.KF;
      for (int row = row0; row < row1; ++row) {
        const int b = row / s.To, to = row % s.To;
        const TX* xr = x + (((long)b * s.Ti + (to * s.st + kt)) * s.Fi + kf) * s.Ci + ci;
        const float* dyr = dy + ((long)row * s.Fo) * s.Co + co;
        for (int fo = 0; fo < s.Fo; ++fo) acc = fmaf(ldf(xr + (long)fo * s.sf * s.Ci), dyr[(long)fo * s.Co], acc);
      }
      atomicAdd(dw + i, acc);
    } else if (dbias) {
      const int co = i - nw;
      for (int row = row0; row < row1; ++row)
        for (int fo = 0; fo < s.Fo; ++fo) acc += dy[((long)row * s.Fo + fo) * s.Co + co];
      atomicAdd(dbias + co, acc);
    }
  }
}

// ------------------------------------------------------------------ batch-stat BN + clipped ReLU, channel-last [R,C]
// The "activation side" tensor (output of fwd / upstream grad of bwd) is addressed through
// element strides (sb,st,sf,sc) over (b,t,f,c) so it can live in NHWC or in the RNN layout.
struct ActMap {
  int T, F, C;          // R = B*T*F rows of C channels
  long sb, st, sf, sc;
};
__device__ __forceinline__ long act_index(const ActMap& m, long r, int c) {
  if (r < 2147483647L) {                      // 32-bit divisions (64-bit ones cost ~100 instructions each)
    const unsigned int ru = (unsigned int)r;
    const unsigned int q = ru / (unsigned int)m.F;
    const unsigned int f = ru - q * (unsigned int)m.F;
    const unsigned int b = q / (unsigned int)m.T;
    const unsigned int t = q - b * (unsigned int)m.T;
    return (long)b * m.sb + (long)t * m.st + (long)f * m.sf + (long)c * m.sc;
  }
  const int f = r % m.F;
  const long q = r / m.F;
  const int t = q % m.T;
  const long b = q / m.T;
  return b * m.sb + t * m.st + f * m.sf + (long)c * m.sc;
}

__global__ void __launch_bounds__(256)
bn_stats_kernel(const float* __restrict__ x, long R, int C, double* __restrict__ sums) {
  // thread -> fixed channel (blockDim multiple of C is required: C | 256)
  const int c = threadIdx.x % C;
  const int rl = threadIdx.x / C, rstep = blockDim.x / C;
  float s1 = 0.f, s2 = 0.f;
  const long rows_per_block = (R + gridDim.x - 1) / gridDim.x;
  const long r0 = blockIdx.x * rows_per_block, r1 = min(R, r0 + rows_per_block);
  for (long r = r0 + rl; r < r1; r += rstep) {
    const float v = x[r * C + c];
    s1 += v;
    s2 = fmaf(v, v, s2);
  }
  __shared__ float sh1[256], sh2[256];
  sh1[threadIdx.x] = s1;
  sh2[threadIdx.x] = s2;
  __syncthreads();
  if (threadIdx.x < C) {
    double a = 0, b = 0;
    for (int i = threadIdx.x; i < blockDim.x; i += C) { a += sh1[i]; b += sh2[i]; }
    atomicAdd(sums + c, a);
    atomicAdd(sums + C + c, b);
  }
}

__global__ void bn_finalize_kernel(const double* __restrict__ sums, double count, float eps, int C, float* __restrict__ mean,
                                   float* __restrict__ rstd) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double m = sums[c] / count;
  double var = sums[C + c] / count - m * m;
  if (var < 0) var = 0;
  mean[c] = (float)m;
  rstd[c] = (float)(1.0 / sqrt(var + (double)eps));
}

template <typename TO>
__global__ void __launch_bounds__(256)
bn_clip_fwd_kernel(const float* __restrict__ x, long R, const float* __restrict__ mean, const float* __restrict__ rstd,
                   const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip, TO* __restrict__ out,
                   ActMap m) {
  // thread <-> fixed channel (blockDim is a multiple of C), rows strided over the grid
  const int C = m.C;
  const int c = threadIdx.x % C;
  const int rl = threadIdx.x / C, rstep = blockDim.x / C;
  const float a = rstd[c] * gamma[c];
  const float sh = beta[c] - mean[c] * a;
  for (long r = (long)blockIdx.x * rstep + rl; r < R; r += (long)gridDim.x * rstep) {
    float v = fmaf(x[r * C + c], a, sh);
    if (clip) v = fminf(fmaxf(v, 0.f), cap);
    stf(out + act_index(m, r, c), v);
  }
}

// pass 1 of the backward: per-channel sum(dbn) and sum(dbn * xhat), dbn = dout * [0 < bn < cap]
template <typename TG>
__global__ void __launch_bounds__(256)
bn_clip_bwd_reduce_kernel(const float* __restrict__ x, long R, const float* __restrict__ mean, const float* __restrict__ rstd,
                          const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip,
                          const TG* __restrict__ dout, ActMap m, double* __restrict__ sums) {
  const int C = m.C;
  const int c = threadIdx.x % C;
  const int rl = threadIdx.x / C, rstep = blockDim.x / C;
  const float mu = mean[c], rs = rstd[c], g = gamma[c], be = beta[c];
  float s1 = 0.f, s2 = 0.f;
  const long rows_per_block = (R + gridDim.x - 1) / gridDim.x;
  const long r0 = blockIdx.x * rows_per_block, r1 = min(R, r0 + rows_per_block);
  for (long r = r0 + rl; r < r1; r += rstep) {
    const float xh = (x[r * C + c] - mu) * rs;
    const float bn = xh * g + be;
    float d = ldf(dout + act_index(m, r, c));
    if (clip && !(bn > 0.f && bn < cap)) d = 0.f;
    s1 += d;
    s2 = fmaf(d, xh, s2);
  }
  __shared__ float sh1[256], sh2[256];
  sh1[threadIdx.x] = s1;
  sh2[threadIdx.x] = s2;
  __syncthreads();
  if (threadIdx.x < C) {
    double a = 0, b = 0;
    for (int i = threadIdx.x; i < blockDim.x; i += C) { a += sh1[i]; b += sh2[i]; }
    atomicAdd(sums + c, a);
    atomicAdd(sums + C + c, b);
  }
}

// pass 2: dx = gamma*rstd*(dbn - mean(dbn) - xhat*mean(dbn*xhat)); dgamma = sum(dbn*xhat); dbeta = sum(dbn)
template <typename TG>
__global__ void __launch_bounds__(256)
bn_clip_bwd_apply_kernel(const float* __restrict__ x, long R, const float* __restrict__ mean, const float* __restrict__ rstd,
                         const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip,
                         const TG* __restrict__ dout, ActMap m, const double* __restrict__ sums, float* __restrict__ dx,
                         float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int C = m.C;
  const int c = threadIdx.x % C;
  const int rl = threadIdx.x / C, rstep = blockDim.x / C;
  const float mu = mean[c], rs = rstd[c], g = gamma[c], be = beta[c];
  const float m1 = (float)(sums[c] / (double)R), m2 = (float)(sums[C + c] / (double)R);
  const float gr = g * rs;
  for (long r = (long)blockIdx.x * rstep + rl; r < R; r += (long)gridDim.x * rstep) {
    const float xh = (x[r * C + c] - mu) * rs;
    const float bn = fmaf(xh, g, be);
    float d = ldf(dout + act_index(m, r, c));
    if (clip && !(bn > 0.f && bn < cap)) d = 0.f;
    dx[r * C + c] = gr * (d - m1 - xh * m2);
  }
  if (blockIdx.x == 0 && threadIdx.x < C) {
    if (dbeta) dbeta[threadIdx.x] = (float)sums[threadIdx.x];
    if (dgamma) dgamma[threadIdx.x] = (float)sums[C + threadIdx.x];
  }
}

// ---- the same three passes for the plain channel-last case (activation side contiguous like x, C % 4 == 0):
// four channels per thread, 16-byte accesses, no index arithmetic.  Thread <-> fixed channel quad, rows strided.
template <typename T> __device__ __forceinline__ float4 ld4(const T* p);
template <> __device__ __forceinline__ float4 ld4<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
template <> __device__ __forceinline__ float4 ld4<__nv_bfloat16>(const __nv_bfloat16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x), b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
  return make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st4(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}
struct Quad { float v[4]; };
__device__ __forceinline__ Quad q4(float4 f) { Quad q; q.v[0] = f.x; q.v[1] = f.y; q.v[2] = f.z; q.v[3] = f.w; return q; }

template <typename TO>
__global__ void __launch_bounds__(256)
bn_clip_fwd_vec_kernel(const float* __restrict__ x, long R, int C, const float* __restrict__ mean, const float* __restrict__ rstd,
                       const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip, TO* __restrict__ out) {
  const int Q = C / 4;
  const int c = (threadIdx.x % Q) * 4;
  const int rl = threadIdx.x / Q, rstep = blockDim.x / Q;
  float a[4], sh[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { a[i] = rstd[c + i] * gamma[c + i]; sh[i] = beta[c + i] - mean[c + i] * a[i]; }
  for (long r = (long)blockIdx.x * rstep + rl; r < R; r += (long)gridDim.x * rstep) {
    const Quad xv = q4(ld4(x + r * C + c));
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      o[i] = fmaf(xv.v[i], a[i], sh[i]);
      if (clip) o[i] = fminf(fmaxf(o[i], 0.f), cap);
    }
    st4(out + r * C + c, make_float4(o[0], o[1], o[2], o[3]));
  }
}

template <typename TG>
__global__ void __launch_bounds__(256)
bn_clip_bwd_reduce_vec_kernel(const float* __restrict__ x, long R, int C, const float* __restrict__ mean, const float* __restrict__ rstd,
                              const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip,
                              const TG* __restrict__ dout, double* __restrict__ sums) {
  const int Q = C / 4;
  const int c = (threadIdx.x % Q) * 4;
  const int rl = threadIdx.x / Q, rstep = blockDim.x / Q;
  float mu[4], rs[4], g[4], be[4], s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 4; ++i) { mu[i] = mean[c + i]; rs[i] = rstd[c + i]; g[i] = gamma[c + i]; be[i] = beta[c + i]; }
  const long rows_per_block = (R + gridDim.x - 1) / gridDim.x;
  const long r0 = blockIdx.x * rows_per_block, r1 = min(R, r0 + rows_per_block);
#pragma unroll 4
  for (long r = r0 + rl; r < r1; r += rstep) {
    const Quad xv = q4(ld4(x + r * C + c)), dv = q4(ld4(dout + r * C + c));
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float xh = (xv.v[i] - mu[i]) * rs[i];
      const float bn = fmaf(xh, g[i], be[i]);
      const float d = (clip && !(bn > 0.f && bn < cap)) ? 0.f : dv.v[i];
      s1[i] += d;
      s2[i] = fmaf(d, xh, s2[i]);
    }
  }
  __shared__ float sh1[4][256], sh2[4][256];
#pragma unroll
  for (int i = 0; i < 4; ++i) { sh1[i][threadIdx.x] = s1[i]; sh2[i][threadIdx.x] = s2[i]; }
  __syncthreads();
  if (threadIdx.x < C) {
    // channel threadIdx.x: quad q = ch / 4, component i = ch % 4, held by threads q, q + Q, q + 2Q, ...
    const int q = threadIdx.x / 4, i = threadIdx.x % 4;
    double a = 0, b = 0;
    for (int t = q; t < blockDim.x; t += Q) { a += sh1[i][t]; b += sh2[i][t]; }
    atomicAdd(sums + threadIdx.x, a);
    atomicAdd(sums + C + threadIdx.x, b);
  }
}

template <typename TG>
__global__ void __launch_bounds__(256)
bn_clip_bwd_apply_vec_kernel(const float* __restrict__ x, long R, int C, const float* __restrict__ mean, const float* __restrict__ rstd,
                             const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip,
                             const TG* __restrict__ dout, const double* __restrict__ sums, float* __restrict__ dx,
                             float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int Q = C / 4;
  const int c = (threadIdx.x % Q) * 4;
  const int rl = threadIdx.x / Q, rstep = blockDim.x / Q;
  float mu[4], rs[4], g[4], be[4], m1[4], m2[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    mu[i] = mean[c + i]; rs[i] = rstd[c + i]; g[i] = gamma[c + i]; be[i] = beta[c + i];
    m1[i] = (float)(sums[c + i] / (double)R);
    m2[i] = (float)(sums[C + c + i] / (double)R);
  }
  for (long r = (long)blockIdx.x * rstep + rl; r < R; r += (long)gridDim.x * rstep) {
    const Quad xv = q4(ld4(x + r * C + c)), dv = q4(ld4(dout + r * C + c));
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float xh = (xv.v[i] - mu[i]) * rs[i];
      const float bn = fmaf(xh, g[i], be[i]);
      const float d = (clip && !(bn > 0.f && bn < cap)) ? 0.f : dv.v[i];
      o[i] = g[i] * rs[i] * (d - m1[i] - xh * m2[i]);
    }
    st4(dx + r * C + c, make_float4(o[0], o[1], o[2], o[3]));
  }
  if (blockIdx.x == 0 && threadIdx.x < C) {
    if (dbeta) dbeta[threadIdx.x] = (float)sums[threadIdx.x];
    if (dgamma) dgamma[threadIdx.x] = (float)sums[C + threadIdx.x];
  }
}

// ---- the three passes when the activation side is the RECURRENT layout [T,B,feat]: one (b,t) row of F*C values at a
// time goes through shared memory, so that both sides are read / written contiguously.  feat = c*F+f (NCHW model,
// src/deepSpeech_NCHW.py:166-168: the row is transposed on the way) or f*C+c (NHWC model, src/deepSpeech.py:162-164).
// Row r = b*T + t of x lives at out row t*B + b.  Thread <-> fixed channel (256 % C == 0); smem holds F*C floats.
struct RowMap {
  int B, T, F, C;
  int cf;              // 1: feat = c*F + f, 0: feat = f*C + c
};
__device__ __forceinline__ long rnn_row(const RowMap& m, int r) {
  const int b = r / m.T, t = r - b * m.T;
  return ((long)t * m.B + b) * ((long)m.F * m.C);
}
__device__ __forceinline__ int feat_slot(const RowMap& m, int f, int c) { return m.cf ? c * m.F + f : f * m.C + c; }

template <typename TO>
__global__ void __launch_bounds__(256)
bn_clip_fwd_rows_kernel(const float* __restrict__ x, int R, const float* __restrict__ mean, const float* __restrict__ rstd,
                        const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip, TO* __restrict__ out,
                        RowMap m) {
  extern __shared__ float bn_row[];
  const int C = m.C, FC = m.F * m.C;
  const int c = threadIdx.x % C, fl = threadIdx.x / C, fstep = blockDim.x / C;
  const float a = rstd[c] * gamma[c];
  const float sh = beta[c] - mean[c] * a;
  for (int r = blockIdx.x; r < R; r += gridDim.x) {
    const float* xr = x + (long)r * FC;
    for (int f = fl; f < m.F; f += fstep) {
      float v = fmaf(xr[f * C + c], a, sh);
      if (clip) v = fminf(fmaxf(v, 0.f), cap);
      bn_row[feat_slot(m, f, c)] = v;
    }
    __syncthreads();
    TO* orow = out + rnn_row(m, r);
    if (sizeof(TO) == 2 && (FC & 7) == 0) {
      // bf16 output: 8 values = one 16-byte store (rows are 16-byte aligned when FC % 8 == 0)
      for (int j = threadIdx.x * 8; j < FC; j += blockDim.x * 8) {
        const float4 a = *reinterpret_cast<const float4*>(bn_row + j), b = *reinterpret_cast<const float4*>(bn_row + j + 4);
        __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w);
        __nv_bfloat162 p2 = __floats2bfloat162_rn(b.x, b.y), p3 = __floats2bfloat162_rn(b.z, b.w);
        uint4 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&p0); pk.y = *reinterpret_cast<uint32_t*>(&p1);
        pk.z = *reinterpret_cast<uint32_t*>(&p2); pk.w = *reinterpret_cast<uint32_t*>(&p3);
        *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(orow) + j) = pk;
      }
    } else {
      for (int j = threadIdx.x; j < FC; j += blockDim.x) stf(orow + j, bn_row[j]);
    }
    __syncthreads();
  }
}

template <typename TG, bool APPLY>
__global__ void __launch_bounds__(256)
bn_clip_bwd_rows_kernel(const float* __restrict__ x, int R, const float* __restrict__ mean, const float* __restrict__ rstd,
                        const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip,
                        const TG* __restrict__ dout, RowMap m, double* __restrict__ sums, float* __restrict__ dx,
                        float* __restrict__ dgamma, float* __restrict__ dbeta) {
  extern __shared__ float bn_row[];
  const int C = m.C, FC = m.F * m.C;
  const int c = threadIdx.x % C, fl = threadIdx.x / C, fstep = blockDim.x / C;
  const float mu = mean[c], rs = rstd[c], g = gamma[c], be = beta[c];
  float m1 = 0.f, m2 = 0.f;
  if (APPLY) {
    m1 = (float)(sums[c] / ((double)R * m.F));
    m2 = (float)(sums[C + c] / ((double)R * m.F));
  }
  float s1 = 0.f, s2 = 0.f;
  for (int r = blockIdx.x; r < R; r += gridDim.x) {
    const TG* drow = dout + rnn_row(m, r);
    if (sizeof(TG) == 2 && (FC & 7) == 0) {
      for (int j = threadIdx.x * 8; j < FC; j += blockDim.x * 8) {
        const uint4 pk = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(drow) + j);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&pk);
        *reinterpret_cast<float4*>(bn_row + j) = make_float4(__low2float(h[0]), __high2float(h[0]), __low2float(h[1]), __high2float(h[1]));
        *reinterpret_cast<float4*>(bn_row + j + 4) = make_float4(__low2float(h[2]), __high2float(h[2]), __low2float(h[3]), __high2float(h[3]));
      }
    } else if (sizeof(TG) == 4 && (FC & 3) == 0) {
      for (int j = threadIdx.x * 4; j < FC; j += blockDim.x * 4)
        *reinterpret_cast<float4*>(bn_row + j) = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(drow) + j);
    } else {
      for (int j = threadIdx.x; j < FC; j += blockDim.x) bn_row[j] = ldf(drow + j);
    }
    __syncthreads();
    const float* xr = x + (long)r * FC;
    for (int f = fl; f < m.F; f += fstep) {
      const float xh = (xr[f * C + c] - mu) * rs;
      const float bn = fmaf(xh, g, be);
      float d = bn_row[feat_slot(m, f, c)];
      if (clip && !(bn > 0.f && bn < cap)) d = 0.f;
      if (APPLY) {
        dx[(long)r * FC + f * C + c] = g * rs * (d - m1 - xh * m2);
      } else {
        s1 += d;
        s2 = fmaf(d, xh, s2);
      }
    }
    __syncthreads();
  }
  if (!APPLY) {
    __shared__ float sh1[256], sh2[256];
    sh1[threadIdx.x] = s1;
    sh2[threadIdx.x] = s2;
    __syncthreads();
    if (threadIdx.x < C) {
      double a = 0, b = 0;
      for (int i = threadIdx.x; i < blockDim.x; i += C) { a += sh1[i]; b += sh2[i]; }
      atomicAdd(sums + c, a);
      atomicAdd(sums + C + c, b);
    }
  } else if (blockIdx.x == 0 && threadIdx.x < C) {
    if (dbeta) dbeta[threadIdx.x] = (float)sums[threadIdx.x];
    if (dgamma) dgamma[threadIdx.x] = (float)sums[C + threadIdx.x];
  }
}

// ------------------------------------------------------------------ element-wise helpers
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) cast_kernel(const TI* __restrict__ in, TO* __restrict__ out, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) stf(out + i, ldf(in + i));
}
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) add2_kernel(const TI* __restrict__ a, const TI* __restrict__ b, TO* __restrict__ out, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    stf(out + i, ldf(a + i) + ldf(b + i));
}
__global__ void __launch_bounds__(256) scale_kernel(const float* __restrict__ x, float s, float* __restrict__ y, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) y[i] = s * x[i];
}
__global__ void __launch_bounds__(256) relux_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long n, float cap, int has_cap) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    float v = fmaxf(x[i], 0.f);
    if (has_cap) v = fminf(v, cap);
    y[i] = v;
  }
}
__global__ void __launch_bounds__(256) relux_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx, long n, float cap, int has_cap) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const float v = x[i];
    dx[i] = (v > 0.f && (!has_cap || v < cap)) ? dy[i] : 0.f;
  }
}
// out[c] (+)= sum_r x[r, c]   (x is [R, N] with row pitch ld)
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, long R, int N, long ld, float* __restrict__ out) {
  const int n = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rl = threadIdx.x >> 5;
  float s = 0.f;
  if (n < N) {
    const long rows_per_block = (R + gridDim.y - 1) / gridDim.y;
    const long r0 = blockIdx.y * rows_per_block, r1 = min(R, r0 + rows_per_block);
    for (long r = r0 + rl; r < r1; r += 8) s += ldf(x + r * ld + n);
  }
  __shared__ float sh[8][33];
  sh[rl][threadIdx.x & 31] = s;
  __syncthreads();
  if (rl == 0 && n < N) {
    float a = 0.f;
    for (int i = 0; i < 8; ++i) a += sh[i][threadIdx.x];
    atomicAdd(out + n, a);
  }
}
// 2-D transpose with dtype conversion: out[c, r] = in[r, c]
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) transpose_kernel(const TI* __restrict__ in, long R, long C, long ldi, TO* __restrict__ out, long ldo) {
  __shared__ float tile[32][33];
  const long r0 = blockIdx.y * 32L, c0 = blockIdx.x * 32L;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int i = ty; i < 32; i += 8)
    if (r0 + i < R && c0 + tx < C) tile[i][tx] = ldf(in + (r0 + i) * ldi + c0 + tx);
  __syncthreads();
  for (int i = ty; i < 32; i += 8)
    if (c0 + i < C && r0 + tx < R) stf(out + (c0 + i) * ldo + r0 + tx, tile[tx][i]);
}

static inline int ew_grid(long n) {
  long g = (n + 255) / 256;
  long cap = (long)num_sms() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace ds2

using namespace ds2;

// ====================================================================== C ABI
extern "C" int ds2_gemm_generic(const void* A, int a_dtype, long sam, long sak, const void* B, int b_dtype, long sbn, long sbk,
                                void* C, int c_dtype, long ldc, int M, int N, int K, float alpha, float beta, const float* bias,
                                void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DS2_REQUIRE(A && B && C, DS2_ERR_BAD_ARG, "ds2_gemm_generic: null pointer");
  DS2_REQUIRE(M >= 0 && N >= 0 && K >= 0, DS2_ERR_BAD_SHAPE, "ds2_gemm_generic: negative dimension");
  if (M == 0 || N == 0) return DS2_OK;
  const int key = a_dtype * 4 + b_dtype * 2 + c_dtype;
  switch (key) {
    case 0: return gemm_generic_launch<float, float, float>(A, sam, sak, B, sbn, sbk, C, ldc, M, N, K, alpha, beta, bias, st);
    case 1: return gemm_generic_launch<float, float, __nv_bfloat16>(A, sam, sak, B, sbn, sbk, C, ldc, M, N, K, alpha, beta, bias, st);
    case 2: return gemm_generic_launch<float, __nv_bfloat16, float>(A, sam, sak, B, sbn, sbk, C, ldc, M, N, K, alpha, beta, bias, st);
    case 4: return gemm_generic_launch<__nv_bfloat16, float, float>(A, sam, sak, B, sbn, sbk, C, ldc, M, N, K, alpha, beta, bias, st);
    case 6: return gemm_generic_launch<__nv_bfloat16, __nv_bfloat16, float>(A, sam, sak, B, sbn, sbk, C, ldc, M, N, K, alpha, beta, bias, st);
    case 7: return gemm_generic_launch<__nv_bfloat16, __nv_bfloat16, __nv_bfloat16>(A, sam, sak, B, sbn, sbk, C, ldc, M, N, K, alpha, beta, bias, st);
    default: break;
  }
  set_error("ds2_gemm_generic: unsupported dtype combination a=%d b=%d c=%d", a_dtype, b_dtype, c_dtype);
  return DS2_ERR_BAD_ARG;
}

static int conv_shape(ConvShape* s, int B, int Ti, int Fi, int Ci, int Co, int KT, int KF, int st, int sf) {
  DS2_REQUIRE(B > 0 && Ti >= KT && Fi >= KF && Ci > 0 && Co > 0 && KT > 0 && KF > 0 && st > 0 && sf > 0, DS2_ERR_BAD_SHAPE,
              "ds2_conv2d: bad shape B=%d Ti=%d Fi=%d Ci=%d Co=%d K=%dx%d stride=%dx%d", B, Ti, Fi, Ci, Co, KT, KF, st, sf);
  s->B = B; s->Ti = Ti; s->Fi = Fi; s->Ci = Ci; s->Co = Co; s->KT = KT; s->KF = KF; s->st = st; s->sf = sf;
  s->To = (Ti - KT) / st + 1;
  s->Fo = (Fi - KF) / sf + 1;
  return DS2_OK;
}

extern "C" int ds2_conv2d_fwd(const void* x, int x_dtype, const float* w, const float* bias, float* y, int B, int Ti, int Fi,
                              int Ci, int Co, int KT, int KF, int st, int sf, void* stream) {
  ConvShape s;
  int rc = conv_shape(&s, B, Ti, Fi, Ci, Co, KT, KF, st, sf);
  if (rc) return rc;
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  const long total = (long)s.B * s.To * s.Fo * s.Co;
  if (x_dtype == DS2_F32)
    conv_fwd_kernel<float><<<ew_grid(total), 256, 0, cs>>>(static_cast<const float*>(x), w, bias, y, s);
  else
    conv_fwd_kernel<__nv_bfloat16><<<ew_grid(total), 256, 0, cs>>>(static_cast<const __nv_bfloat16*>(x), w, bias, y, s);
  DS2_LAUNCH_CHECK("conv_fwd_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_conv2d_dgrad(const float* dy, const float* w, float* dx, int B, int Ti, int Fi, int Ci, int Co, int KT,
                                int KF, int st, int sf, void* stream) {
  ConvShape s;
  int rc = conv_shape(&s, B, Ti, Fi, Ci, Co, KT, KF, st, sf);
  if (rc) return rc;
  const long total = (long)s.B * s.Ti * s.Fi * s.Ci;
  conv_dgrad_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, w, dx, s);
  DS2_LAUNCH_CHECK("conv_dgrad_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_conv2d_wgrad(const void* x, int x_dtype, const float* dy, float* dw, float* dbias, int B, int Ti, int Fi,
                                int Ci, int Co, int KT, int KF, int st, int sf, void* stream) {
  ConvShape s;
  int rc = conv_shape(&s, B, Ti, Fi, Ci, Co, KT, KF, st, sf);
  if (rc) return rc;
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  const int nw = KT * KF * Ci * Co;
  DS2_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * nw, cs));
  if (dbias) DS2_CUDA(cudaMemsetAsync(dbias, 0, sizeof(float) * Co, cs));
  const int rows = s.B * s.To;
  int rows_per_block = 4;
  dim3 grid((nw + Co + 255) / 256, (rows + rows_per_block - 1) / rows_per_block);
  if (grid.y > 65535) { rows_per_block = (rows + 65534) / 65535; grid.y = (rows + rows_per_block - 1) / rows_per_block; }
  if (x_dtype == DS2_F32)
    conv_wgrad_kernel<float><<<grid, 256, 0, cs>>>(static_cast<const float*>(x), dy, dw, dbias, s, rows_per_block);
  else
    conv_wgrad_kernel<__nv_bfloat16><<<grid, 256, 0, cs>>>(static_cast<const __nv_bfloat16*>(x), dy, dw, dbias, s, rows_per_block);
  DS2_LAUNCH_CHECK("conv_wgrad_kernel");
  count_launch();
  return DS2_OK;
}

static int act_map(ActMap* m, int T, int F, int C, long sb, long st, long sf, long sc) {
  DS2_REQUIRE(C > 0 && C <= 256, DS2_ERR_BAD_SHAPE, "ds2_bn: channel count %d must be in [1,256]", C);
  m->T = T; m->F = F; m->C = C; m->sb = sb; m->st = st; m->sf = sf; m->sc = sc;
  return DS2_OK;
}

// activation side laid out exactly like x ([B,T,F,C] contiguous)?
static bool act_contiguous(const ActMap& m, int B) {
  (void)B;
  return m.sc == 1 && m.sf == m.C && m.st == (long)m.F * m.C && m.sb == (long)m.T * m.F * m.C;
}
static bool vec_channels(int C) { return C % 4 == 0 && 256 % (C / 4) == 0; }
// activation side = recurrent layout [T,B,F*C] with feat = c*F+f (cf) or f*C+c ?
static bool act_rnn_rows(const ActMap& m, int B, RowMap* rm) {
  const long feat = (long)m.F * m.C;
  if (256 % m.C != 0 || feat * 4 > 40 * 1024 || m.sb != feat || m.st != (long)B * feat) return false;
  if (m.sf == 1 && m.sc == m.F) rm->cf = 1;
  else if (m.sc == 1 && m.sf == m.C) rm->cf = 0;
  else return false;
  rm->B = B; rm->T = m.T; rm->F = m.F; rm->C = m.C;
  return true;
}

extern "C" int ds2_bn_stats(const float* x, long R, int C, float eps, double* sums_scratch, float* mean, float* rstd, void* stream) {
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  DS2_REQUIRE(x && sums_scratch && mean && rstd && R > 0, DS2_ERR_BAD_ARG, "ds2_bn_stats: bad argument");
  DS2_REQUIRE(C > 0 && C <= 256, DS2_ERR_BAD_SHAPE, "ds2_bn: channel count %d must be in [1,256]", C);
  DS2_CUDA(cudaMemsetAsync(sums_scratch, 0, sizeof(double) * 2 * C, cs));
  long rows_per = 256 / C * 16;
  int grid = (int)((R + rows_per - 1) / rows_per);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  bn_stats_kernel<<<grid, (256 / C) * C, 0, cs>>>(x, R, C, sums_scratch);
  DS2_LAUNCH_CHECK("bn_stats_kernel");
  bn_finalize_kernel<<<1, 256, 0, cs>>>(sums_scratch, (double)R, eps, C, mean, rstd);
  DS2_LAUNCH_CHECK("bn_finalize_kernel");
  count_launch(2);
  return DS2_OK;
}

// mean / rstd from per-channel sums a producer kernel accumulated (sums[c] = sum x, sums[C + c] = sum x^2 over `count` values)
extern "C" int ds2_bn_finalize(const double* sums, double count, int C, float eps, float* mean, float* rstd, void* stream) {
  DS2_REQUIRE(sums && mean && rstd && count > 0 && C > 0 && C <= 256, DS2_ERR_BAD_ARG, "ds2_bn_finalize: bad argument");
  bn_finalize_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(sums, count, eps, C, mean, rstd);
  DS2_LAUNCH_CHECK("bn_finalize_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_bn_clip_fwd(const float* x, int B, int T, int F, int C, const float* mean, const float* rstd,
                               const float* gamma, const float* beta, float cap, int clip, void* out, int out_dtype, long sb,
                               long st, long sf, long sc, void* stream) {
  ActMap m;
  int rc = act_map(&m, T, F, C, sb, st, sf, sc);
  if (rc) return rc;
  const long R = (long)B * T * F;
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (act_contiguous(m, B) && vec_channels(C)) {
    const int threads = (256 / (C / 4)) * (C / 4);
    if (out_dtype == DS2_F32)
      bn_clip_fwd_vec_kernel<float><<<ew_grid(R * C / 4), threads, 0, cs>>>(x, R, C, mean, rstd, gamma, beta, cap, clip, static_cast<float*>(out));
    else
      bn_clip_fwd_vec_kernel<__nv_bfloat16><<<ew_grid(R * C / 4), threads, 0, cs>>>(x, R, C, mean, rstd, gamma, beta, cap, clip, static_cast<__nv_bfloat16*>(out));
    DS2_LAUNCH_CHECK("bn_clip_fwd_vec_kernel");
    count_launch();
    return DS2_OK;
  }
  RowMap rm;
  if (R < 2147483647L / C && act_rnn_rows(m, B, &rm)) {
    const int rows = (int)((long)B * T);
    const int grid = rows < num_sms() * 8 ? rows : num_sms() * 8;
    const size_t sm = (size_t)F * C * sizeof(float);
    if (out_dtype == DS2_F32)
      bn_clip_fwd_rows_kernel<float><<<grid, 256, sm, cs>>>(x, rows, mean, rstd, gamma, beta, cap, clip, static_cast<float*>(out), rm);
    else
      bn_clip_fwd_rows_kernel<__nv_bfloat16><<<grid, 256, sm, cs>>>(x, rows, mean, rstd, gamma, beta, cap, clip, static_cast<__nv_bfloat16*>(out), rm);
    DS2_LAUNCH_CHECK("bn_clip_fwd_rows_kernel");
    count_launch();
    return DS2_OK;
  }
  if (out_dtype == DS2_F32)
    bn_clip_fwd_kernel<float><<<ew_grid(R * C), (256 / C) * C, 0, cs>>>(x, R, mean, rstd, gamma, beta, cap, clip, static_cast<float*>(out), m);
  else
    bn_clip_fwd_kernel<__nv_bfloat16><<<ew_grid(R * C), (256 / C) * C, 0, cs>>>(x, R, mean, rstd, gamma, beta, cap, clip, static_cast<__nv_bfloat16*>(out), m);
  DS2_LAUNCH_CHECK("bn_clip_fwd_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_bn_clip_bwd(const float* x, int B, int T, int F, int C, const float* mean, const float* rstd,
                               const float* gamma, const float* beta, float cap, int clip, const void* dout, int dout_dtype,
                               long sb, long st, long sf, long sc, double* sums_scratch, float* dx, float* dgamma, float* dbeta,
                               void* stream) {
  ActMap m;
  int rc = act_map(&m, T, F, C, sb, st, sf, sc);
  if (rc) return rc;
  const long R = (long)B * T * F;
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  DS2_CUDA(cudaMemsetAsync(sums_scratch, 0, sizeof(double) * 2 * C, cs));
  long rows_per = 256 / C * 16;
  int grid = (int)((R + rows_per - 1) / rows_per);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  if (act_contiguous(m, B) && vec_channels(C)) {
    const int threads = (256 / (C / 4)) * (C / 4);
    if (dout_dtype == DS2_F32) {
      const float* d = static_cast<const float*>(dout);
      bn_clip_bwd_reduce_vec_kernel<float><<<grid, threads, 0, cs>>>(x, R, C, mean, rstd, gamma, beta, cap, clip, d, sums_scratch);
      DS2_LAUNCH_CHECK("bn_clip_bwd_reduce_vec_kernel");
      bn_clip_bwd_apply_vec_kernel<float><<<ew_grid(R * C / 4), threads, 0, cs>>>(x, R, C, mean, rstd, gamma, beta, cap, clip, d, sums_scratch, dx, dgamma, dbeta);
    } else {
      const __nv_bfloat16* d = static_cast<const __nv_bfloat16*>(dout);
      bn_clip_bwd_reduce_vec_kernel<__nv_bfloat16><<<grid, threads, 0, cs>>>(x, R, C, mean, rstd, gamma, beta, cap, clip, d, sums_scratch);
      DS2_LAUNCH_CHECK("bn_clip_bwd_reduce_vec_kernel");
      bn_clip_bwd_apply_vec_kernel<__nv_bfloat16><<<ew_grid(R * C / 4), threads, 0, cs>>>(x, R, C, mean, rstd, gamma, beta, cap, clip, d, sums_scratch, dx, dgamma, dbeta);
    }
    DS2_LAUNCH_CHECK("bn_clip_bwd_apply_vec_kernel");
    count_launch(2);
    return DS2_OK;
  }
  RowMap rm;
  if (R < 2147483647L / C && act_rnn_rows(m, B, &rm)) {
    const int rows = (int)((long)B * T);
    const int g2 = rows < num_sms() * 8 ? rows : num_sms() * 8;
    const size_t sm = (size_t)F * C * sizeof(float);
    if (dout_dtype == DS2_F32) {
      const float* d = static_cast<const float*>(dout);
      bn_clip_bwd_rows_kernel<float, false><<<g2, 256, sm, cs>>>(x, rows, mean, rstd, gamma, beta, cap, clip, d, rm, sums_scratch, dx, dgamma, dbeta);
      DS2_LAUNCH_CHECK("bn_clip_bwd_rows_kernel<reduce>");
      bn_clip_bwd_rows_kernel<float, true><<<g2, 256, sm, cs>>>(x, rows, mean, rstd, gamma, beta, cap, clip, d, rm, sums_scratch, dx, dgamma, dbeta);
    } else {
      const __nv_bfloat16* d = static_cast<const __nv_bfloat16*>(dout);
      bn_clip_bwd_rows_kernel<__nv_bfloat16, false><<<g2, 256, sm, cs>>>(x, rows, mean, rstd, gamma, beta, cap, clip, d, rm, sums_scratch, dx, dgamma, dbeta);
      DS2_LAUNCH_CHECK("bn_clip_bwd_rows_kernel<reduce>");
      bn_clip_bwd_rows_kernel<__nv_bfloat16, true><<<g2, 256, sm, cs>>>(x, rows, mean, rstd, gamma, beta, cap, clip, d, rm, sums_scratch, dx, dgamma, dbeta);
    }
    DS2_LAUNCH_CHECK("bn_clip_bwd_rows_kernel<apply>");
    count_launch(2);
    return DS2_OK;
  }
  if (dout_dtype == DS2_F32) {
    const float* d = static_cast<const float*>(dout);
    bn_clip_bwd_reduce_kernel<float><<<grid, (256 / C) * C, 0, cs>>>(x, R, mean, rstd, gamma, beta, cap, clip, d, m, sums_scratch);
    DS2_LAUNCH_CHECK("bn_clip_bwd_reduce_kernel");
    bn_clip_bwd_apply_kernel<float><<<ew_grid(R * C), (256 / C) * C, 0, cs>>>(x, R, mean, rstd, gamma, beta, cap, clip, d, m, sums_scratch, dx, dgamma, dbeta);
  } else {
    const __nv_bfloat16* d = static_cast<const __nv_bfloat16*>(dout);
    bn_clip_bwd_reduce_kernel<__nv_bfloat16><<<grid, (256 / C) * C, 0, cs>>>(x, R, mean, rstd, gamma, beta, cap, clip, d, m, sums_scratch);
    DS2_LAUNCH_CHECK("bn_clip_bwd_reduce_kernel");
    bn_clip_bwd_apply_kernel<__nv_bfloat16><<<ew_grid(R * C), (256 / C) * C, 0, cs>>>(x, R, mean, rstd, gamma, beta, cap, clip, d, m, sums_scratch, dx, dgamma, dbeta);
  }
  DS2_LAUNCH_CHECK("bn_clip_bwd_apply_kernel");
  count_launch(2);
  return DS2_OK;
}

extern "C" int ds2_cast(const void* in, int in_dtype, void* out, int out_dtype, long n, void* stream) {
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (n <= 0) return DS2_OK;
  if (in_dtype == DS2_F32 && out_dtype == DS2_BF16)
    cast_kernel<float, __nv_bfloat16><<<ew_grid(n), 256, 0, cs>>>(static_cast<const float*>(in), static_cast<__nv_bfloat16*>(out), n);
  else if (in_dtype == DS2_BF16 && out_dtype == DS2_F32)
    cast_kernel<__nv_bfloat16, float><<<ew_grid(n), 256, 0, cs>>>(static_cast<const __nv_bfloat16*>(in), static_cast<float*>(out), n);
  else {
    set_error("ds2_cast: unsupported dtype pair %d -> %d", in_dtype, out_dtype);
    return DS2_ERR_BAD_ARG;
  }
  DS2_LAUNCH_CHECK("cast_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_add2(const void* a, const void* b, int in_dtype, void* out, int out_dtype, long n, void* stream) {
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (n <= 0) return DS2_OK;
  const int key = in_dtype * 2 + out_dtype;
  if (key == 0) add2_kernel<float, float><<<ew_grid(n), 256, 0, cs>>>((const float*)a, (const float*)b, (float*)out, n);
  else if (key == 1) add2_kernel<float, __nv_bfloat16><<<ew_grid(n), 256, 0, cs>>>((const float*)a, (const float*)b, (__nv_bfloat16*)out, n);
  else if (key == 2) add2_kernel<__nv_bfloat16, float><<<ew_grid(n), 256, 0, cs>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (float*)out, n);
  else add2_kernel<__nv_bfloat16, __nv_bfloat16><<<ew_grid(n), 256, 0, cs>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)out, n);
  DS2_LAUNCH_CHECK("add2_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_scale(const float* x, float s, float* y, long n, void* stream) {
  if (n <= 0) return DS2_OK;
  scale_kernel<<<ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, s, y, n);
  DS2_LAUNCH_CHECK("scale_kernel");
  count_launch();
  return DS2_OK;
}
extern "C" int ds2_relux_fwd(const float* x, float* y, long n, float cap, int has_cap, void* stream) {
  if (n <= 0) return DS2_OK;
  relux_fwd_kernel<<<ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, n, cap, has_cap);
  DS2_LAUNCH_CHECK("relux_fwd_kernel");
  count_launch();
  return DS2_OK;
}
extern "C" int ds2_relux_bwd(const float* x, const float* dy, float* dx, long n, float cap, int has_cap, void* stream) {
  if (n <= 0) return DS2_OK;
  relux_bwd_kernel<<<ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, dy, dx, n, cap, has_cap);
  DS2_LAUNCH_CHECK("relux_bwd_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_colsum(const void* x, int dtype, long R, int N, long ld, float* out, int accumulate, void* stream) {
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (!accumulate) DS2_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * N, cs));
  if (R <= 0 || N <= 0) return DS2_OK;
  int gy = (int)((R + 511) / 512);
  if (gy > 1024) gy = 1024;
  dim3 grid((N + 31) / 32, gy);
  if (dtype == DS2_F32) colsum_kernel<float><<<grid, 256, 0, cs>>>((const float*)x, R, N, ld, out);
  else colsum_kernel<__nv_bfloat16><<<grid, 256, 0, cs>>>((const __nv_bfloat16*)x, R, N, ld, out);
  DS2_LAUNCH_CHECK("colsum_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_transpose(const void* in, int in_dtype, long R, long C, long ldi, void* out, int out_dtype, long ldo, void* stream) {
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (R <= 0 || C <= 0) return DS2_OK;
  dim3 grid((unsigned)((C + 31) / 32), (unsigned)((R + 31) / 32));
  const int key = in_dtype * 2 + out_dtype;
  if (key == 0) transpose_kernel<float, float><<<grid, 256, 0, cs>>>((const float*)in, R, C, ldi, (float*)out, ldo);
  else if (key == 1) transpose_kernel<float, __nv_bfloat16><<<grid, 256, 0, cs>>>((const float*)in, R, C, ldi, (__nv_bfloat16*)out, ldo);
  else if (key == 2) transpose_kernel<__nv_bfloat16, float><<<grid, 256, 0, cs>>>((const __nv_bfloat16*)in, R, C, ldi, (float*)out, ldo);
  else transpose_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, cs>>>((const __nv_bfloat16*)in, R, C, ldi, (__nv_bfloat16*)out, ldo);
  DS2_LAUNCH_CHECK("transpose_kernel");
  count_launch();
  return DS2_OK;
}
