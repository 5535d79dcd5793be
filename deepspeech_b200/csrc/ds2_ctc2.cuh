// CTC forward-backward, second form (round 2); included by ds2_ctc.cu.  Replaces tf.nn.ctc_loss as called at
// src/deepSpeech_NCHW.py:225 / src/deepSpeech.py:227 (blank = K-1, softmax inside, gradient w.r.t. the logits).
//
// The round-1 chain kernel kept one lattice state per thread in log space: every step was LDS x4 -> EX2 x3 -> LG2 ->
// STS -> CTA barrier, 340-520 ns per step.  Here the recursion runs in the LINEAR domain with scaling (Rabiner):
//   * a fully parallel first kernel turns every logits row into yhat = exp(x - max_k x) (largest class = 1) and keeps
//     log Z = log sum_k yhat: no transcendental is left on the T-step dependency chain;
//   * one WARP per (utterance, direction): lane l keeps states [l*NS, (l+1)*NS) in registers; a step is
//       v(s) = a(s) + a(s-1) + skip(s) * a(s-2);   a(s) <- v(s) * yhat_t(class(s))
//     with the one neighbour below a lane's range fetched by a warp shuffle - no shared-memory hand-off, no CTA
//     barrier;
//   * every CTC2_RB steps the warp rescales by a power of two (exponent of the largest state, integer bookkeeping:
//     exact), so values stay inside fp32's range however long the utterance is;
//   * table rows are staged in shared memory and leave through one bulk copy per block; the emission rows of the next
//     block are fetched while this one is computed;
//   * a third, fully parallel kernel forms softmax - occupancy per (t, b) row from the two tables.
// alpha stores the value with the step's emission, beta without it, so that alpha_t(s) * beta_t(s) * 2^(Ea + Eb - Efin)
// / p_hat is the posterior of state s at time t (the products of the Z_t cancel).
// A row whose label cannot be aligned (or whose probability underflows fp32 entirely) has p_hat = 0: its loss is +inf
// and its gradient row is zero - tf.nn.ctc_loss raises for such rows; callers that cannot check on the host see the inf.
#pragma once

namespace ds2 {
namespace c2 {

static constexpr int CTC2_RB = 8;          // steps per block (rescale + bulk copy granularity)
static constexpr int CTC2_KP = 32;         // class slots per emission row (K <= 32)

struct Ws2 {
  float* alpha;      // [B][T][pitch]
  float* beta;       // [B][T][pitch]
  float* yhat;       // [T][B][32]    exp(x - rowmax), 0 beyond K
  float* lz;         // [T][B]        log sum_k yhat
  int* ea;           // [B][NB]       cumulative power-of-two exponent applied before block j (alpha), NB = blocks + 1
  int* eb;           // [B][NB]       same for beta (block index counts steps from the END of the utterance)
  float* phat;       // [B]           alpha_hat_{Tb-1}(S-1) + alpha_hat_{Tb-1}(S-2)
  int* efin;         // [B]           exponent that goes with phat
  int* perm;         // [B][Lmax]     label positions sorted by class
  int* cls_off;      // [B][K+1]
  int pitch, NB, NS;
};

__host__ __device__ inline int ctc2_ns(int Lmax) {
  const int S = 2 * Lmax + 1;
  int ns = (S + 31) / 32;
  ns += ns & 1;                    // even: a state's parity is then the parity of its register index
  return ns < 2 ? 2 : ns;
}

static size_t ctc2_ws_layout(int T, int B, int K, int Lmax, char* base, Ws2* ws) {
  const int NS = ctc2_ns(Lmax), pitch = 32 * NS;
  const int NB = (T + CTC2_RB - 1) / CTC2_RB + 1;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~size_t(255); return o; };
  const size_t o_a = take(sizeof(float) * (size_t)B * T * pitch);
  const size_t o_b = take(sizeof(float) * (size_t)B * T * pitch);
  const size_t o_y = take(sizeof(float) * (size_t)T * B * CTC2_KP);
  const size_t o_lz = take(sizeof(float) * (size_t)T * B);
  const size_t o_ea = take(sizeof(int) * (size_t)B * NB);
  const size_t o_eb = take(sizeof(int) * (size_t)B * NB);
  const size_t o_ph = take(sizeof(float) * (size_t)B);
  const size_t o_ef = take(sizeof(int) * (size_t)B);
  const size_t o_perm = take(sizeof(int) * (size_t)B * (Lmax > 0 ? Lmax : 1));
  const size_t o_cls = take(sizeof(int) * (size_t)B * (K + 1));
  if (ws) {
    ws->alpha = reinterpret_cast<float*>(base + o_a);
    ws->beta = reinterpret_cast<float*>(base + o_b);
    ws->yhat = reinterpret_cast<float*>(base + o_y);
    ws->lz = reinterpret_cast<float*>(base + o_lz);
    ws->ea = reinterpret_cast<int*>(base + o_ea);
    ws->eb = reinterpret_cast<int*>(base + o_eb);
    ws->phat = reinterpret_cast<float*>(base + o_ph);
    ws->efin = reinterpret_cast<int*>(base + o_ef);
    ws->perm = reinterpret_cast<int*>(base + o_perm);
    ws->cls_off = reinterpret_cast<int*>(base + o_cls);
    ws->pitch = pitch;
    ws->NB = NB;
    ws->NS = NS;
  }
  return off;
}

// ---- kernel 1: one warp per logits row (t, b): yhat = exp(x - max), log Z
__global__ void __launch_bounds__(256) ctc2_rows_kernel(const float* __restrict__ logits, long rows, int K, float* __restrict__ yhat,
                                                        float* __restrict__ lz) {
  const int lane = threadIdx.x & 31;
  const long row = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  const float x = lane < K ? __ldg(logits + row * K + lane) : -INFINITY;
  const float m = warp_max(x);
  const float e = lane < K ? __expf(x - m) : 0.f;
  const float z = warp_sum(e);
  yhat[row * CTC2_KP + lane] = e;
  if (lane == 0) lz[row] = __logf(z);
}

__device__ __forceinline__ void c2_bulk_store_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void c2_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void c2_bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void c2_bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---- kernel 2: the two recursions, one warp each.  DIR 0: alpha, forward in time; DIR 1: beta = the same recursion on
// the reversed label sequence, backward in time.  Tables are written in the chain's own state order (beta: reversed).
template <int DIR, int NS>
__device__ __forceinline__ void ctc2_chain(int T, int B, int K, const int* __restrict__ labels, int Lmax,
                                           const int* __restrict__ label_lens, const int* __restrict__ seq_lens,
                                           float* __restrict__ loss, const Ws2& ws, float* smem) {
  const int b = blockIdx.x, lane = threadIdx.x;
  const int pitch = 32 * NS;
  float* em = smem;                                  // [2][RB][32] emission rows of a block
  float* stage = smem + 2 * CTC2_RB * CTC2_KP;       // [2][RB][pitch] table rows of a block
  int* lab = reinterpret_cast<int*>(stage + 2 * CTC2_RB * pitch);   // [Lmax] (reversed for DIR 1)
  const int Tb = min(max(seq_lens[b], 0), T);
  int L = label_lens[b];
  const bool bad = L < 0 || L > Lmax;                // the caller's label row is unusable: treated as not alignable
  if (bad) L = 0;
  const int S = 2 * L + 1;
  const int blank = K - 1;
  float* tab = (DIR == 0 ? ws.alpha : ws.beta) + (size_t)b * T * pitch;
  int* ecum = (DIR == 0 ? ws.ea : ws.eb) + (size_t)b * ws.NB;

  for (int i = lane; i < L; i += 32) lab[i] = labels[(size_t)b * Lmax + (DIR == 0 ? i : L - 1 - i)];
  __syncwarp();
  if (DIR == 0) {
    // class-sorted (stable) permutation of label positions for the gradient kernel: lane k owns class k
    int* cls = ws.cls_off + (size_t)b * (K + 1);
    int* perm = ws.perm + (size_t)b * Lmax;
    int cnt = 0;
    for (int i = 0; i < L; ++i) cnt += (lab[i] == lane);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int c = incl - cnt;
    if (lane < K) cls[lane] = c;
    for (int i = 0; i < L; ++i)
      if (lab[i] == lane) perm[c++] = i;
    if (lane == 31) cls[K] = incl;
  }
  if (Tb <= 0) {
    if (DIR == 0 && lane == 0) { loss[b] = bad ? INFINITY : 0.f; ws.phat[b] = bad ? 0.f : 1.f; ws.efin[b] = 0; }
    return;
  }
  // sum_t log Z_t (fp64, off the chain)
  double lzsum = 0.0;
  if (DIR == 0) {
    for (int t = lane; t < Tb; t += 32) lzsum += (double)ws.lz[(size_t)t * B + b];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lzsum += __shfl_xor_sync(0xffffffffu, lzsum, o);
  }
  // per-lane lattice description: state s = lane*NS + j; even j = blank, odd j = label (s >> 1)
  const int s0 = lane * NS;
  const int limit = min(max(S - s0, 0), NS);        // live states of this lane
  int ycls[NS / 2];
  uint32_t skipmask = 0u;
#pragma unroll
  for (int h = 0; h < NS / 2; ++h) {
    const int s = s0 + 2 * h + 1;
    const int li = s >> 1;
    ycls[h] = (s < S) ? lab[li] : blank;
    if (s < S && li >= 1 && lab[li] != lab[li - 1]) skipmask |= 1u << h;
  }
  float a[NS];
#pragma unroll
  for (int j = 0; j < NS; ++j) a[j] = 0.f;

  // emission rows of block 0
  {
    const int n0 = min(CTC2_RB, Tb);
    for (int r = 0; r < n0; ++r) {
      const int t = (DIR == 0) ? r : (Tb - 1 - r);
      em[r * CTC2_KP + lane] = ws.yhat[((size_t)t * B + b) * CTC2_KP + lane];
    }
  }
  __syncwarp();
  int cum = 0;
  if (lane == 0) ecum[0] = 0;
  const int nblk = (Tb + CTC2_RB - 1) / CTC2_RB;
  for (int blk = 0; blk < nblk; ++blk) {
    const int step0 = blk * CTC2_RB;
    const int nsteps = min(CTC2_RB, Tb - step0);
    const int nnext = min(CTC2_RB, Tb - step0 - CTC2_RB);
    const float* emc = em + (blk & 1) * CTC2_RB * CTC2_KP;
    float* emn = em + ((blk + 1) & 1) * CTC2_RB * CTC2_KP;
    float* stg = stage + (blk & 1) * CTC2_RB * pitch;
    // the next block's emission rows -> registers now, shared memory at the end of this block
    float pre[CTC2_RB];
#pragma unroll
    for (int r = 0; r < CTC2_RB; ++r) {
      pre[r] = 0.f;
      if (r < nnext) {
        const int st = step0 + CTC2_RB + r;
        const int t = (DIR == 0) ? st : (Tb - 1 - st);
        pre[r] = __ldg(ws.yhat + ((size_t)t * B + b) * CTC2_KP + lane);
      }
    }
    for (int i = 0; i < nsteps; ++i) {
      const float* row = emc + i * CTC2_KP;
      const float yb = row[blank];
      float yl[NS / 2];
#pragma unroll
      for (int h = 0; h < NS / 2; ++h) yl[h] = row[ycls[h]];
      // the only value a lane needs from its neighbour: the last state of the lane below (state s0 - 1).  State s0 is a
      // blank (NS is even) and has no skip transition, so s0 - 2 is never read.
      float p1 = __shfl_up_sync(0xffffffffu, a[NS - 1], 1);
      if (lane == 0) p1 = 0.f;
      float v[NS];
      if (step0 + i == 0) {
#pragma unroll
        for (int j = 0; j < NS; ++j) v[j] = (s0 + j < 2) ? 1.f : 0.f;
      } else {
#pragma unroll
        for (int j = NS - 1; j >= 0; --j) {
          const float am1 = (j >= 1) ? a[j - 1] : p1;
          float x = a[j] + am1;
          if (j & 1) {                                  // label state: may skip the blank below it
            const float am2 = (j >= 2) ? a[j - 2] : p1;          // j == 1: s - 2 is the last state of the lane below
            if (skipmask & (1u << (j >> 1))) x += am2;
          }
          v[j] = x;
        }
      }
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        const float y = (j & 1) ? yl[j >> 1] : yb;
        const bool live = j < limit;
        v[j] = live ? v[j] : 0.f;
        a[j] = live ? v[j] * y : 0.f;
      }
      // table row: alpha with the emission, beta without
      float* dst = stg + ((DIR == 0) ? i : (nsteps - 1 - i)) * pitch + s0;
#pragma unroll
      for (int j = 0; j < NS; j += 2)
        *reinterpret_cast<float2*>(dst + j) = (DIR == 0) ? make_float2(a[j], a[j + 1]) : make_float2(v[j], v[j + 1]);
    }
    // ---- block boundary
    const bool last = (step0 + nsteps == Tb);
    if (!last) {
      // rescale by the power of two of the largest state (exact; the exponent is remembered per block)
      float m = 0.f;
#pragma unroll
      for (int j = 0; j < NS; ++j) m = fmaxf(m, a[j]);
      m = warp_max(m);
      const int ex = (int)((__float_as_uint(m) >> 23) & 0xFFu);
      const int e = ex == 0 ? 0 : ex - 127;                      // m == 0 (or flushed): nothing to scale, p_hat will be 0
      const float sc = __uint_as_float((uint32_t)(127 - e) << 23);
#pragma unroll
      for (int j = 0; j < NS; ++j) a[j] *= sc;
      cum += e;
      if (lane == 0) ecum[blk + 1] = cum;
#pragma unroll
      for (int r = 0; r < CTC2_RB; ++r)
        if (r < nnext) emn[r * CTC2_KP + lane] = pre[r];
    } else if (DIR == 0) {
      // p_hat = alpha_hat(S-1) + alpha_hat(S-2) of the last step, read back from the staged row
      __syncwarp();
      if (lane == 0) {
        const float* lastrow = stg + (nsteps - 1) * pitch;
        const float ph = bad ? 0.f : lastrow[S - 1] + (S > 1 ? lastrow[S - 2] : 0.f);
        ws.phat[b] = ph;
        ws.efin[b] = cum;
        loss[b] = (ph > 0.f) ? (float)(-((double)__logf(ph) + (double)cum * 0.6931471805599453 - lzsum)) : INFINITY;
      }
    }
    fence_proxy_async();                 // staged rows (generic proxy) -> visible to the bulk copy (async proxy)
    __syncwarp();
    if (lane == 0) {
      const int t_lo = (DIR == 0) ? step0 : (Tb - step0 - nsteps);
      c2_bulk_store_s2g(tab + (size_t)t_lo * pitch, stg, (uint32_t)(nsteps * pitch * 4));
      c2_bulk_commit();
      c2_bulk_wait_read1();              // the previous block's copy has finished reading its staging buffer
    }
    __syncwarp();
  }
  if (lane == 0) c2_bulk_wait_all();
}

template <int NS>
__global__ void __launch_bounds__(32) ctc2_chain_kernel(int T, int B, int K, const int* __restrict__ labels, int Lmax,
                                                        const int* __restrict__ label_lens, const int* __restrict__ seq_lens,
                                                        float* __restrict__ loss, Ws2 ws) {
  extern __shared__ __align__(16) float c2_smem[];
  if (blockIdx.y == 0) ctc2_chain<0, NS>(T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws, c2_smem);
  else ctc2_chain<1, NS>(T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws, c2_smem);
}

// ---- kernel 3: one warp per (t, b) row: grad = (softmax - occupancy) * dloss.  dynamic smem per warp: float e[pitch]
__global__ void __launch_bounds__(256) ctc2_grad_kernel(int T, int B, int K, int Lmax, const int* __restrict__ label_lens,
                                                        const int* __restrict__ seq_lens, const float* __restrict__ dloss,
                                                        float* __restrict__ grad, Ws2 ws) {
  extern __shared__ __align__(16) float c2_gsm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int b = blockIdx.y;
  const int t = blockIdx.x * nwarps + warp;
  if (t >= T) return;
  float* e = c2_gsm + (size_t)warp * ws.pitch;
  const int Tb = min(max(seq_lens[b], 0), T);
  int L = label_lens[b];
  if (L < 0 || L > Lmax) L = 0;
  const int S = 2 * L + 1;
  const int blank = K - 1;
  float* g = grad + ((size_t)t * B + b) * K;
  const float ph = ws.phat[b];
  if (t >= Tb || !(ph > 0.f)) {                       // outside the utterance, or a row that cannot be aligned
    if (lane < K) g[lane] = 0.f;
    return;
  }
  const float up = dloss ? dloss[b] : 1.f;
  const int ex = ws.ea[(size_t)b * ws.NB + t / CTC2_RB] + ws.eb[(size_t)b * ws.NB + (Tb - 1 - t) / CTC2_RB] - ws.efin[b];
  // 2^ex / p_hat, split over the two factors so that neither product leaves fp32's range on the way
  const int e1 = max(-125, min(125, ex / 2)), e2 = max(-125, min(125, ex - ex / 2));
  const float f1 = __uint_as_float((uint32_t)(e1 + 127) << 23), f2 = __uint_as_float((uint32_t)(e2 + 127) << 23) / ph;
  const float* ar = ws.alpha + ((size_t)b * T + t) * ws.pitch;
  const float* br = ws.beta + ((size_t)b * T + t) * ws.pitch;     // beta is stored in reversed state order
  float blank_sum = 0.f;
  for (int s = lane; s < S; s += 32) {
    const float v = fminf((ar[s] * f1) * (br[S - 1 - s] * f2), 1.0009765625f);
    e[s] = v;
    if (!(s & 1)) blank_sum += v;
  }
  blank_sum = warp_sum(blank_sum);
  __syncwarp();
  const int* perm = ws.perm + (size_t)b * Lmax;
  const int* cls = ws.cls_off + (size_t)b * (K + 1);
  if (lane < K) {
    const float y = ws.yhat[((size_t)t * B + b) * CTC2_KP + lane] * __expf(-ws.lz[(size_t)t * B + b]);
    float occ;
    if (lane == blank) {
      occ = blank_sum;
    } else {
      occ = 0.f;
      for (int i = cls[lane]; i < cls[lane + 1]; ++i) occ += e[2 * perm[i] + 1];
    }
    g[lane] = (y - occ) * up;
  }
}

}  // namespace c2
}  // namespace ds2
