// CTC forward-backward, second form (round 2); included by ds2_ctc.cu.  Replaces tf.nn.ctc_loss as called at
// src/deepSpeech_NCHW.py:225 / src/deepSpeech.py:227 (blank = K-1, softmax inside, gradient w.r.t. the logits).
//
// The round-1 chain kernel kept one lattice state per thread and exchanged neighbours through shared memory: every
// step was LDS x4 -> EX2 x3 -> LG2 -> STS x2 -> CTA barrier over up to 26 warps, 340-520 ns per step.  Here:
//   * a fully parallel first kernel turns every logits row into its log2-softmax, so no row reduction and no
//     conversion is left on the T-step dependency chain;
//   * a chain (utterance, direction) is run by W <= 4 warps (one per SM sub-partition); lane l of warp w keeps the NS
//     consecutive states [(32w + l) * NS, +NS) in REGISTERS; a step is, in log2 units,
//       v(s) = lse(a(s), a(s-1), skip(s) ? a(s-2) : -inf);   a(s) <- v(s) + ly_t(class(s))
//     with the one neighbour below a lane's range fetched by a warp shuffle (across warps: one float through shared
//     memory); NS is even, so a lane's first state is a blank and never needs s-2 from below;
//   * lse costs 2 MUFU (blank) or 3 (label): the largest term is 2^0 = 1 and is not exponentiated;
//   * every CTC2_RB steps the block maximum is subtracted (offsets accumulated in fp64), so stored values stay small
//     enough for fp32 to resolve them; table rows are staged in shared memory and leave through one bulk copy per block;
//     the emission rows of the next block are fetched while this one is computed;
//   * a third, fully parallel kernel forms softmax - occupancy per (t, b) row from the two tables.
// alpha stores the value with the step's emission, beta without it: alpha_t(s) + beta_t(s) + offsets - log2 p is the
// log2-posterior of state s at time t.  Tables are in the chain's own state order (beta: reversed).
// Tried first and dropped: the recursion in the LINEAR domain with power-of-two rescaling (no MUFU at all on the chain,
// 206 ns per step at (400,100,64) with one warp per chain).  It is exact in fp64 but not in fp32: the states that carry
// the posterior sit in the far tail of alpha wherever the rest of the utterance forces the alignment forward, more than
// 2^126 below the step's largest state from T = 800 on (emulation against fp64: posterior errors > 1).
// A row whose label cannot be aligned has log2 p = -inf-like: its loss is +inf and its gradient row is zero -
// tf.nn.ctc_loss raises for such rows; callers that cannot check on the host see the inf.
#pragma once

namespace ds2 {
namespace c2 {

static constexpr int CTC2_RB = 16;         // steps per block (renormalisation + bulk copy granularity)
static constexpr float C2_NEG = -1e30f;
static constexpr float C2_LOG2E = 1.4426950408889634f;
static constexpr double C2_LN2 = 0.6931471805599453;
static constexpr int CTC2_KP = 32;         // class slots per emission row (K <= 32)

struct Ws2 {
  float* alpha;      // [B][T][pitch]
  float* beta;       // [B][T][pitch]
  float* ly;         // [T][B][32]    log2-softmax of the logits row, -1e30 beyond K
  double* coff;      // [B][2][NB]    cumulative offset subtracted before block j (alpha, beta), log2 units
  double* logp;      // [B]           log2 p(label | logits)
  int* perm;         // [B][Lmax]     label positions sorted by class
  int* cls_off;      // [B][K+1]
  int pitch, NB, NS, W;
};

// warps per chain and states per lane: up to 4 warps (one per SM sub-partition), NS even (a state's parity is then the
// parity of its register index) and as small as the label allows
__host__ __device__ inline void ctc2_geometry(int Lmax, int* W, int* NS) {
  const int S = 2 * Lmax + 1;
  int w = S <= 64 ? 1 : (S <= 128 ? 2 : 4);
  int ns = (S + 32 * w - 1) / (32 * w);
  ns += ns & 1;
  if (ns < 2) ns = 2;
  *W = w;
  *NS = ns;
}

static size_t ctc2_ws_layout(int T, int B, int K, int Lmax, char* base, Ws2* ws) {
  int W, NS;
  ctc2_geometry(Lmax, &W, &NS);
  const int pitch = (2 * Lmax + 1 + 3) & ~3;        // table row pitch: the states only (the thread grid may be wider)
  const int NB = (T + CTC2_RB - 1) / CTC2_RB + 1;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~size_t(255); return o; };
  const size_t o_a = take(sizeof(float) * (size_t)B * T * pitch);
  const size_t o_b = take(sizeof(float) * (size_t)B * T * pitch);
  const size_t o_y = take(sizeof(float) * (size_t)T * B * CTC2_KP);
  const size_t o_co = take(sizeof(double) * (size_t)B * 2 * NB);
  const size_t o_lp = take(sizeof(double) * (size_t)B);
  const size_t o_perm = take(sizeof(int) * (size_t)B * (Lmax > 0 ? Lmax : 1));
  const size_t o_cls = take(sizeof(int) * (size_t)B * (K + 1));
  if (ws) {
    ws->alpha = reinterpret_cast<float*>(base + o_a);
    ws->beta = reinterpret_cast<float*>(base + o_b);
    ws->ly = reinterpret_cast<float*>(base + o_y);
    ws->coff = reinterpret_cast<double*>(base + o_co);
    ws->logp = reinterpret_cast<double*>(base + o_lp);
    ws->perm = reinterpret_cast<int*>(base + o_perm);
    ws->cls_off = reinterpret_cast<int*>(base + o_cls);
    ws->pitch = pitch;
    ws->NB = NB;
    ws->NS = NS;
    ws->W = W;
  }
  return off;
}

// ---- kernel 1: log2-softmax of every logits row (t, b); a warp takes C2_ROWS rows at a time so that their loads and
// reductions overlap
static constexpr int C2_ROWS = 4;
__global__ void __launch_bounds__(256) ctc2_rows_kernel(const float* __restrict__ logits, long rows, int K, float* __restrict__ ly) {
  const int lane = threadIdx.x & 31;
  const long row0 = ((blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5) * C2_ROWS;
  float x[C2_ROWS];
#pragma unroll
  for (int r = 0; r < C2_ROWS; ++r) x[r] = (lane < K && row0 + r < rows) ? __ldg(logits + (row0 + r) * K + lane) : -INFINITY;
#pragma unroll
  for (int r = 0; r < C2_ROWS; ++r) {
    if (row0 + r >= rows) break;
    const float m = warp_max(x[r]);
    const float e = lane < K ? __expf(x[r] - m) : 0.f;
    const float z = warp_sum(e);
    ly[(row0 + r) * CTC2_KP + lane] = lane < K ? (x[r] - m) * C2_LOG2E - __log2f(z) : C2_NEG;
  }
}

__device__ __forceinline__ void c2_bulk_store_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void c2_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void c2_bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void c2_bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }


// log2(2^a + 2^b) and log2(2^a + 2^b + 2^c): the largest term is 2^0 and costs no MUFU
__device__ __forceinline__ float c2_lse2(float a, float b) {
  const float m = fmaxf(a, b);
  return m + __log2f(1.f + exp2f(fminf(a, b) - m));
}
__device__ __forceinline__ float c2_lse3(float a, float b, float c) {
  const float hi = fmaxf(a, b), lo = fminf(a, b);
  const float m = fmaxf(hi, c);
  return m + __log2f(1.f + exp2f(fminf(hi, c) - m) + exp2f(lo - m));
}

// ---- kernel 2: the two recursions.  DIR 0: alpha, forward in time; DIR 1: beta = the same recursion on the reversed
// label sequence, backward in time.  W warps per chain, NS states per lane.
template <int DIR, int NS, int W>
__device__ __forceinline__ void ctc2_chain(int T, int B, int K, const int* __restrict__ labels, int Lmax,
                                           const int* __restrict__ label_lens, const int* __restrict__ seq_lens,
                                           float* __restrict__ loss, const Ws2& ws, float* smem) {
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NT = 32 * W;
  const int pitch = ws.pitch;                        // row pitch of tables and staging (>= S, multiple of 4)
  float* em = smem;                                  // [2][RB][32] emission rows of a block
  float* stage = em + 2 * CTC2_RB * CTC2_KP;         // [2][RB][pitch] table rows of a block
  float* xch = stage + 2 * CTC2_RB * pitch;          // [2][4] last state of every warp (for the warp above), [8..11] block max
  int* lab = reinterpret_cast<int*>(xch + 16);       // [Lmax] (reversed for DIR 1)
  const int Tb = min(max(seq_lens[b], 0), T);
  int L = label_lens[b];
  const bool bad = L < 0 || L > Lmax;                // the caller's label row is unusable: treated as not alignable
  if (bad) L = 0;
  const int S = 2 * L + 1;
  const int blank = K - 1;
  float* tab = (DIR == 0 ? ws.alpha : ws.beta) + (size_t)b * T * pitch;
  double* coff = ws.coff + ((size_t)b * 2 + DIR) * ws.NB;

  for (int i = tid; i < L; i += NT) lab[i] = labels[(size_t)b * Lmax + (DIR == 0 ? i : L - 1 - i)];
  if (tid < 8) xch[tid] = C2_NEG;
  __syncthreads();
  if (DIR == 0 && warp == 0) {
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
    if (DIR == 0 && tid == 0) { loss[b] = bad ? INFINITY : 0.f; ws.logp[b] = bad ? (double)C2_NEG : 0.0; }
    return;
  }
  // per-thread lattice description: state s = tid*NS + j; even j = blank, odd j = label (s >> 1)
  const int s0 = tid * NS;
  const int limit = min(max(S - s0, 0), NS);        // live states of this thread
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
  for (int j = 0; j < NS; ++j) a[j] = C2_NEG;

  // emission rows of block 0
  {
    const int n0 = min(CTC2_RB, Tb);
    for (int e = tid; e < n0 * CTC2_KP; e += NT) {
      const int r = e >> 5;
      const int t = (DIR == 0) ? r : (Tb - 1 - r);
      em[e] = ws.ly[((size_t)t * B + b) * CTC2_KP + (e & 31)];
    }
  }
  __syncthreads();
  double cum = 0.0;
  if (tid == 0) coff[0] = 0.0;
  int par = 0;
  const int nblk = (Tb + CTC2_RB - 1) / CTC2_RB;
  constexpr int PRE = (CTC2_RB * CTC2_KP + NT - 1) / NT;     // emission values a thread carries for the next block
  for (int blk = 0; blk < nblk; ++blk) {
    const int step0 = blk * CTC2_RB;
    const int nsteps = min(CTC2_RB, Tb - step0);
    const int nnext = min(CTC2_RB, Tb - step0 - CTC2_RB);
    const float* emc = em + (blk & 1) * CTC2_RB * CTC2_KP;
    float* emn = em + ((blk + 1) & 1) * CTC2_RB * CTC2_KP;
    float* stg = stage + (blk & 1) * CTC2_RB * pitch;
    // the next block's emission rows -> registers now, shared memory at the end of this block
    float pre[PRE];
#pragma unroll
    for (int m = 0; m < PRE; ++m) {
      const int e = tid + m * NT;
      pre[m] = 0.f;
      if (e < nnext * CTC2_KP) {
        const int st = step0 + CTC2_RB + (e >> 5);
        const int t = (DIR == 0) ? st : (Tb - 1 - st);
        pre[m] = __ldg(ws.ly + ((size_t)t * B + b) * CTC2_KP + (e & 31));
      }
    }
    for (int i = 0; i < nsteps; ++i) {
      const float* row = emc + i * CTC2_KP;
      const float yb = row[blank];
      float yl[NS / 2];
#pragma unroll
      for (int h = 0; h < NS / 2; ++h) yl[h] = row[ycls[h]];
      // the only value a thread needs from below its range: the last state of the thread below (state s0 - 1).  State
      // s0 is a blank (NS is even) and has no skip transition, so s0 - 2 is never read.
      float p1 = __shfl_up_sync(0xffffffffu, a[NS - 1], 1);
      if (lane == 0) p1 = (W > 1 && warp > 0) ? xch[par * 4 + warp - 1] : C2_NEG;
      float v[NS];
      if (step0 + i == 0) {
#pragma unroll
        for (int j = 0; j < NS; ++j) v[j] = (s0 + j < 2) ? 0.f : C2_NEG;
      } else {
#pragma unroll
        for (int j = 0; j < NS; ++j) {
          const float am1 = (j >= 1) ? a[j - 1] : p1;
          if (j & 1) {                                  // label state: may skip the blank below it
            const float am2 = (j >= 2) ? a[j - 2] : p1;          // j == 1: s - 2 is the last state of the thread below
            v[j] = c2_lse3(a[j], am1, (skipmask & (1u << (j >> 1))) ? am2 : C2_NEG);
          } else {
            v[j] = c2_lse2(a[j], am1);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        const float y = (j & 1) ? yl[j >> 1] : yb;
        const bool live = j < limit;
        v[j] = live ? v[j] : C2_NEG;
        a[j] = live ? v[j] + y : C2_NEG;
      }
      par ^= 1;
      if (W > 1 && lane == 31) xch[par * 4 + warp] = a[NS - 1];
      // table row: alpha with the emission, beta without
      float* dst = stg + ((DIR == 0) ? i : (nsteps - 1 - i)) * pitch + s0;
#pragma unroll
      for (int j = 0; j < NS; j += 2)
        if (s0 + j < pitch) *reinterpret_cast<float2*>(dst + j) = (DIR == 0) ? make_float2(a[j], a[j + 1]) : make_float2(v[j], v[j + 1]);
      if (W > 1) __syncthreads();
    }
    // ---- block boundary
    const bool last = (step0 + nsteps == Tb);
    if (!last) {
      // subtract the block maximum so that stored magnitudes stay small; the offset is remembered per block in fp64
      float m = C2_NEG;
#pragma unroll
      for (int j = 0; j < NS; ++j) m = fmaxf(m, a[j]);
      m = warp_max(m);
      if (W > 1) {
        if (lane == 0) xch[8 + warp] = m;
        __syncthreads();
        m = fmaxf(fmaxf(xch[8], xch[9]), fmaxf(xch[10], xch[11]));
      }
      if (m > 0.5f * C2_NEG) {
#pragma unroll
        for (int j = 0; j < NS; ++j) a[j] = (a[j] > 0.5f * C2_NEG) ? a[j] - m : C2_NEG;
        if (W > 1 && lane == 31) xch[par * 4 + warp] = a[NS - 1];     // the value the warp above reads next step
        cum += (double)m;
      }
      if (tid == 0) coff[blk + 1] = cum;
#pragma unroll
      for (int mm = 0; mm < PRE; ++mm) {
        const int e = tid + mm * NT;
        if (e < nnext * CTC2_KP) emn[e] = pre[mm];
      }
    }
    fence_proxy_async();                 // staged rows (generic proxy) -> visible to the bulk copy (async proxy)
    __syncthreads();
    if (last && DIR == 0 && tid == 0) {
      // log2 p = lse(alpha(S-1), alpha(S-2)) of the last step (read back from the staged row) + the offsets
      const float* lastrow = stg + (nsteps - 1) * pitch;
      const float aL = lastrow[S - 1], aL1 = (S > 1) ? lastrow[S - 2] : C2_NEG;
      const float mx = fmaxf(aL, aL1);
      const bool feasible = !bad && mx > 0.5f * C2_NEG;
      const double lp = feasible ? (double)mx + (double)__log2f(exp2f(aL - mx) + exp2f(aL1 - mx)) + cum : (double)C2_NEG;
      ws.logp[b] = lp;
      loss[b] = feasible ? (float)(-lp * C2_LN2) : INFINITY;
    }
    if (tid == 0) {
      const int t_lo = (DIR == 0) ? step0 : (Tb - step0 - nsteps);
      c2_bulk_store_s2g(tab + (size_t)t_lo * pitch, stg, (uint32_t)(nsteps * pitch * 4));
      c2_bulk_commit();
      c2_bulk_wait_read1();              // the previous block's copy has finished reading its staging buffer
    }
    __syncthreads();
  }
  if (tid == 0) c2_bulk_wait_all();
}

template <int NS, int W>
__global__ void __launch_bounds__(32 * W) ctc2_chain_kernel(int T, int B, int K, const int* __restrict__ labels, int Lmax,
                                                            const int* __restrict__ label_lens, const int* __restrict__ seq_lens,
                                                            float* __restrict__ loss, Ws2 ws) {
  extern __shared__ __align__(16) float c2_smem[];
  if (blockIdx.y == 0) ctc2_chain<0, NS, W>(T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws, c2_smem);
  else ctc2_chain<1, NS, W>(T, B, K, labels, Lmax, label_lens, seq_lens, loss, ws, c2_smem);
}

// ---- kernel 3: grad = (softmax - occupancy) * dloss, one warp per (t, b) row, C2_GROWS rows per warp; the block's rows
// belong to one utterance, whose class-sorted label positions are staged in shared memory once.
// dynamic smem: int perm[Lmax] + int cls[K+1] (rounded up to 4) + per warp float e[pitch]
static constexpr int C2_GROWS = 4;
__global__ void __launch_bounds__(256) ctc2_grad_kernel(int T, int B, int K, int Lmax, const int* __restrict__ label_lens,
                                                        const int* __restrict__ seq_lens, const float* __restrict__ dloss,
                                                        float* __restrict__ grad, Ws2 ws) {
  extern __shared__ __align__(16) float c2_gsm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int b = blockIdx.y;
  int* s_perm = reinterpret_cast<int*>(c2_gsm);
  int* s_cls = s_perm + ((Lmax + 3) & ~3);
  float* e = c2_gsm + ((Lmax + 3) & ~3) + ((K + 1 + 3) & ~3) + (size_t)warp * ws.pitch;
  const int Tb = min(max(seq_lens[b], 0), T);
  int L = label_lens[b];
  if (L < 0 || L > Lmax) L = 0;
  const int S = 2 * L + 1;
  const int blank = K - 1;
  const double logp = ws.logp[b];
  const float up = dloss ? dloss[b] : 1.f;
  for (int i = threadIdx.x; i < L; i += blockDim.x) s_perm[i] = ws.perm[(size_t)b * Lmax + i];
  for (int i = threadIdx.x; i <= K; i += blockDim.x) s_cls[i] = ws.cls_off[(size_t)b * (K + 1) + i];
  __syncthreads();
  const bool feasible = logp > 0.5 * (double)C2_NEG;
  const int t0 = (blockIdx.x * nwarps + warp) * C2_GROWS;
  const int c_lo = lane < K ? s_cls[lane] : 0, c_hi = lane < K ? s_cls[lane + 1] : 0;
  for (int t = t0; t < min(t0 + C2_GROWS, T); ++t) {
    float* g = grad + ((size_t)t * B + b) * K;
    if (t >= Tb || !feasible) {                        // outside the utterance, or a row that cannot be aligned
      if (lane < K) g[lane] = 0.f;
      continue;
    }
    const float yl2 = lane < K ? __ldg(ws.ly + ((size_t)t * B + b) * CTC2_KP + lane) : C2_NEG;
    // alpha rows written at step t carry coff_a[t / RB]; beta rows written at step Tb-1-t carry coff_b[(Tb-1-t) / RB]
    const double ca = ws.coff[((size_t)b * 2 + 0) * ws.NB + t / CTC2_RB];
    const double cb = ws.coff[((size_t)b * 2 + 1) * ws.NB + (Tb - 1 - t) / CTC2_RB];
    const float c = (float)(ca + cb - logp);
    const float* ar = ws.alpha + ((size_t)b * T + t) * ws.pitch;
    const float* br = ws.beta + ((size_t)b * T + t) * ws.pitch;     // beta is stored in reversed state order
    float blank_sum = 0.f;
    for (int s = lane; s < S; s += 32) {
      const float v = exp2f(fminf(__ldg(ar + s) + __ldg(br + S - 1 - s) + c, 1e-3f));
      e[s] = v;
      if (!(s & 1)) blank_sum += v;
    }
    blank_sum = warp_sum(blank_sum);
    __syncwarp();
    if (lane < K) {
      float occ;
      if (lane == blank) {
        occ = blank_sum;
      } else {
        occ = 0.f;
        for (int i = c_lo; i < c_hi; ++i) occ += e[2 * s_perm[i] + 1];
      }
      g[lane] = (exp2f(yl2) - occ) * up;
    }
    __syncwarp();
  }
}

}  // namespace c2
}  // namespace ds2
