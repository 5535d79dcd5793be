// Convolution front-end on the tensor cores: patch matrix (im2col) builders for the tcgen05 GEMM.
//
//   conv fwd : y[M,Co]   = P[M,Kp] . W[Kp,Co]              (ds2_gemm_bf16, B operand MN-major = HWIO filter as is)
//   wgrad    : dW[Kp,Co] = P^T . dy                         (ds2_gemm_bf16_splitk, both operands MN-major)
//   dgrad    : dP[M,Kp]  = dy[M,Co] . W^T  -> ds2_col2im    (gather-sum back to the input grid)
// with M = B*To*Fo rows (b,to,fo) and k = (kt*KF + kf)*Ci + ci, Kp = K rounded up to 8 (zero padded).
// The batch-norm + clipped-ReLU of the PREVIOUS layer (src/custom_ops.py:107-160, :99-104) is applied
// while the patches are gathered, so the normalised activation is never materialised on its own.
// Reference call sites: src/deepSpeech_NCHW.py:120-127, :142-151 (tf.nn.conv2d, VALID, HWIO).
#include "common.cuh"

namespace ds2 {

struct ColGeom {
  int B, Ti, Fi, Ci, To, Fo, KT, KF, st, sf, K;
  long ldp;
};

template <bool BN>
__global__ void __launch_bounds__(256)
im2col_kernel(const float* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ rstd,
              const float* __restrict__ gamma, const float* __restrict__ beta, float cap, int clip,
              __nv_bfloat16* __restrict__ P, ColGeom g) {
  // per-channel affine of the fused batch norm: y = x * sc[c] + sh[c]
  __shared__ __align__(16) float sc[256], sh[256];
  if (BN) {
    for (int c = threadIdx.x; c < g.Ci; c += blockDim.x) {
      const float a = rstd[c] * gamma[c];
      sc[c] = a;
      sh[c] = beta[c] - mean[c] * a;
    }
    __syncthreads();
  }
  // 32-bit index math only: a block works on rows_per_block patch rows at a time, the thread's 8-wide K chunk is fixed
  const int kvec = (int)(g.ldp / 8);
  const int rows_per_block = blockDim.x / kvec > 0 ? blockDim.x / kvec : 1;
  const int M = g.B * g.To * g.Fo;
  const bool vec = (g.Ci % 8) == 0;
  for (int kv0 = 0; kv0 < kvec; kv0 += blockDim.x) {          // more than one pass only when kvec > blockDim
    const int rl = (kvec <= (int)blockDim.x) ? (int)threadIdx.x / kvec : 0;
    const int kv = (kvec <= (int)blockDim.x) ? (int)threadIdx.x - rl * kvec : kv0 + (int)threadIdx.x;
    if (rl >= rows_per_block || kv >= kvec) continue;
    const int k8 = kv * 8;
    // tap decomposition of this thread's chunk (row independent)
    const bool fast = vec && k8 + 8 <= g.K;
    int ci = 0, kf = 0, kt = 0;
    if (fast) {
      ci = k8 % g.Ci;
      const int q = k8 / g.Ci;
      kf = q % g.KF;
      kt = q / g.KF;
    }
    float s8[8], h8[8];
    if (BN && fast) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { s8[i] = sc[ci + i]; h8[i] = sh[ci + i]; }
    }
    // slow path (Ci not a multiple of 8, e.g. the single-channel conv1): the tap of each of the 8 elements is row
    // independent too, so its input offset relative to the patch origin is computed once, not per row
    int off8[8];
    if (!fast) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = k8 + i;
        off8[i] = -1;
        if (k < g.K) {
          const int c1 = k % g.Ci;
          const int q = k / g.Ci;
          const int kf1 = q % g.KF, kt1 = q / g.KF;
          off8[i] = (kt1 * g.Fi + kf1) * g.Ci + c1;
          if (BN) { s8[i] = sc[c1]; h8[i] = sh[c1]; }
        }
      }
    }
    for (int m = blockIdx.x * rows_per_block + rl; m < M; m += gridDim.x * rows_per_block) {
      const int fo = m % g.Fo;
      const int r = m / g.Fo;
      const int to = r % g.To;
      const int b = r / g.To;
      float v[8];
      if (fast) {
        const float* src = x + (((long)b * g.Ti + (to * g.st + kt)) * g.Fi + (fo * g.sf + kf)) * g.Ci + ci;
        const float4 a = __ldg(reinterpret_cast<const float4*>(src));
        const float4 c = __ldg(reinterpret_cast<const float4*>(src) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
        if (BN) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float t = fmaf(v[i], s8[i], h8[i]);
            v[i] = clip ? fminf(fmaxf(t, 0.f), cap) : t;
          }
        }
      } else {
        const float* src = x + (((long)b * g.Ti + to * g.st) * g.Fi + fo * g.sf) * g.Ci;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float t = 0.f;
          if (off8[i] >= 0) {
            t = __ldg(src + off8[i]);
            if (BN) {
              t = fmaf(t, s8[i], h8[i]);
              if (clip) t = fminf(fmaxf(t, 0.f), cap);
            }
          }
          v[i] = t;
        }
      }
      uint4 pk;
      __nv_bfloat162 b0 = __floats2bfloat162_rn(v[0], v[1]), b1 = __floats2bfloat162_rn(v[2], v[3]);
      __nv_bfloat162 b2 = __floats2bfloat162_rn(v[4], v[5]), b3 = __floats2bfloat162_rn(v[6], v[7]);
      pk.x = *reinterpret_cast<uint32_t*>(&b0); pk.y = *reinterpret_cast<uint32_t*>(&b1);
      pk.z = *reinterpret_cast<uint32_t*>(&b2); pk.w = *reinterpret_cast<uint32_t*>(&b3);
      *reinterpret_cast<uint4*>(P + (long)m * g.ldp + k8) = pk;
    }
  }
}

// dx[b,ti,fi,ci..ci+7] = sum over taps (kt,kf) whose output position exists of dP[m(b,to,fo)][(kt*KF+kf)*Ci + ci..]
__global__ void __launch_bounds__(256)
col2im_kernel(const __nv_bfloat16* __restrict__ dP, float* __restrict__ dx, ColGeom g) {
  const int cvec = g.Ci / 8;
  const int total = g.B * g.Ti * g.Fi * cvec;              // < 2^31 (checked by the host wrapper): 32-bit index math
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int ci = (idx % cvec) * 8;
    int r = idx / cvec;
    const int fi = r % g.Fi; r /= g.Fi;
    const int ti = r % g.Ti;
    const int b = r / g.Ti;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int kt = 0; kt < g.KT; ++kt) {
      const int tn = ti - kt;
      if (tn < 0 || (tn % g.st)) continue;
      const int to = tn / g.st;
      if (to >= g.To) continue;
      for (int kf = 0; kf < g.KF; ++kf) {
        const int fn = fi - kf;
        if (fn < 0 || (fn % g.sf)) continue;
        const int fo = fn / g.sf;
        if (fo >= g.Fo) continue;
        const long m = ((long)b * g.To + to) * g.Fo + fo;
        const uint4 pk = __ldg(reinterpret_cast<const uint4*>(dP + m * g.ldp + (kt * g.KF + kf) * g.Ci + ci));
        const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&pk);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += __bfloat162float(h[i]);
      }
    }
    float4* dst = reinterpret_cast<float4*>(dx + (((long)b * g.Ti + ti) * g.Fi + fi) * g.Ci + ci);
    dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}

static int col_geom(ColGeom* g, int B, int Ti, int Fi, int Ci, int KT, int KF, int st, int sf, long ldp) {
  DS2_REQUIRE(B > 0 && Ti >= KT && Fi >= KF && Ci > 0 && KT > 0 && KF > 0 && st > 0 && sf > 0, DS2_ERR_BAD_SHAPE,
              "ds2_im2col: bad shape B=%d Ti=%d Fi=%d Ci=%d K=%dx%d stride=%dx%d", B, Ti, Fi, Ci, KT, KF, st, sf);
  g->B = B; g->Ti = Ti; g->Fi = Fi; g->Ci = Ci; g->KT = KT; g->KF = KF; g->st = st; g->sf = sf;
  g->To = (Ti - KT) / st + 1;
  g->Fo = (Fi - KF) / sf + 1;
  g->K = KT * KF * Ci;
  g->ldp = ldp;
  DS2_REQUIRE(ldp >= g->K && ldp % 8 == 0, DS2_ERR_BAD_SHAPE, "ds2_im2col: row pitch %ld must be a multiple of 8 and >= K=%d", ldp, g->K);
  return DS2_OK;
}

static inline int grid_for(long n) {
  long gsz = (n + 255) / 256;
  const long cap = (long)num_sms() * 32;
  return (int)(gsz < 1 ? 1 : (gsz > cap ? cap : gsz));
}

}  // namespace ds2

using namespace ds2;

extern "C" int ds2_im2col(const float* x, const float* mean, const float* rstd, const float* gamma, const float* beta, float cap,
                          int clip, void* P, long ldp, int B, int Ti, int Fi, int Ci, int KT, int KF, int st, int sf,
                          void* stream) {
  ColGeom g;
  int rc = col_geom(&g, B, Ti, Fi, Ci, KT, KF, st, sf, ldp);
  if (rc) return rc;
  DS2_REQUIRE(x && P, DS2_ERR_BAD_ARG, "ds2_im2col: null pointer");
  DS2_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(P) & 15) == 0, DS2_ERR_BAD_ARG,
              "ds2_im2col: pointers must be 16-byte aligned");
  DS2_REQUIRE(Ci <= 256, DS2_ERR_BAD_SHAPE, "ds2_im2col: at most 256 input channels (got %d)", Ci);
  DS2_REQUIRE((long)g.B * g.To * g.Fo < 2147483647L, DS2_ERR_BAD_SHAPE, "ds2_im2col: too many patch rows");
  const int kvec_h = (int)(ldp / 8);
  const int rpb = 256 / kvec_h > 0 ? 256 / kvec_h : 1;
  const long total = ((long)g.B * g.To * g.Fo + rpb - 1) / rpb * 256;      // one block-iteration per rows_per_block rows
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (mean) {
    DS2_REQUIRE(rstd && gamma && beta, DS2_ERR_BAD_ARG, "ds2_im2col: incomplete batch-norm parameters");
    im2col_kernel<true><<<grid_for(total), 256, 0, cs>>>(x, mean, rstd, gamma, beta, cap, clip, static_cast<__nv_bfloat16*>(P), g);
  } else {
    im2col_kernel<false><<<grid_for(total), 256, 0, cs>>>(x, nullptr, nullptr, nullptr, nullptr, cap, 0, static_cast<__nv_bfloat16*>(P), g);
  }
  DS2_LAUNCH_CHECK("im2col_kernel");
  count_launch();
  return DS2_OK;
}

extern "C" int ds2_col2im(const void* dP, long ldp, float* dx, int B, int Ti, int Fi, int Ci, int KT, int KF, int st, int sf,
                          void* stream) {
  ColGeom g;
  int rc = col_geom(&g, B, Ti, Fi, Ci, KT, KF, st, sf, ldp);
  if (rc) return rc;
  DS2_REQUIRE(dP && dx, DS2_ERR_BAD_ARG, "ds2_col2im: null pointer");
  DS2_REQUIRE(Ci % 8 == 0, DS2_ERR_UNSUPPORTED, "ds2_col2im: input channels (%d) must be a multiple of 8", Ci);
  const long total = (long)g.B * g.Ti * g.Fi * (g.Ci / 8);
  DS2_REQUIRE(total < 2147483647L, DS2_ERR_BAD_SHAPE, "ds2_col2im: tensor too large for 32-bit indexing");
  col2im_kernel<<<grid_for(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(dP), dx, g);
  DS2_LAUNCH_CHECK("col2im_kernel");
  count_launch();
  return DS2_OK;
}
