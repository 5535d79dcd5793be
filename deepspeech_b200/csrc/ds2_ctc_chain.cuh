// CTC alpha / beta recursion (one CTA per (utterance, direction)); included by ds2_ctc.cu.
//
// The loop is a T-step dependency chain with one CTA barrier per step, so nothing slow may be outstanding
// at a barrier: bar.sync waits until the participating threads' prior global accesses are performed, and a
// per-step STG (table row) or LDG (emission) costs an L2 round trip (~600 clk) per step (measured: 340 ns /
// step with them, see profiles/).  Therefore, per block of CTC_RB = 16 steps:
//   * emissions of the NEXT block are fetched into registers at the top of the block and parked in shared
//     memory at its end (one exposed global latency per 16 steps);
//   * table rows are staged in shared memory and written out by ONE asynchronous bulk copy per block
//     (cp.async.bulk shared -> global, async proxy: never waited on by bar.sync), double buffered;
//   * the re-normalisation (block max subtracted, offset accumulated in fp64) happens at the block boundary and
//     its offsets stay in shared memory until the end.
// Values are in log2 units (MUFU.EX2 / MUFU.LG2 without conversion multiplies on the chain).
#pragma once

namespace ds2 {

static constexpr int CTC_RB = RENORM;     // steps per block
static constexpr int CTC_PRE = 8;         // next-block emission values held in registers per thread

struct CtcSmem {
  int o_buf0, o_buf1, o_lse, o_lab, o_red, o_coffs, o_em, o_stage;   // offsets in floats
  int kp, staged;
  size_t bytes;
};
__host__ __device__ inline CtcSmem ctc_smem_layout(int T, int Lmax, int K, int Sp) {
  CtcSmem m;
  const int Smax = 2 * Lmax + 1;
  const int a = (Smax + 2 + 3) & ~3;
  const int nb = (T + CTC_RB - 1) / CTC_RB + 1;
  int off = 0;
  m.o_buf0 = off; off += a;
  m.o_buf1 = off; off += a;
  m.o_lse = off; off += (T + 3) & ~3;
  m.o_lab = off; off += ((Lmax > 0 ? Lmax : 1) + 3) & ~3;
  m.o_red = off; off += 32;
  m.o_coffs = off; off += 2 * ((nb + 1) & ~1);          // doubles (offset is a multiple of 4 floats)
  m.kp = (K + 3) & ~3;
  m.o_em = off; off += 2 * CTC_RB * m.kp;
  m.o_stage = off;
  const size_t stage_floats = 2 * (size_t)CTC_RB * Sp;
  m.staged = ((size_t)off + stage_floats) * 4 <= 190 * 1024 ? 1 : 0;
  m.bytes = ((size_t)off + (m.staged ? stage_floats : 0)) * 4;
  return m;
}

__device__ __forceinline__ void bulk_store_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

template <int DIR, int NS>
__device__ void ctc_chain(const float* __restrict__ logits, int T, int B, int K, const int* __restrict__ labels, int Lmax,
                          const int* __restrict__ label_lens, const int* __restrict__ seq_lens, float* __restrict__ loss,
                          CtcWs ws, float* smem) {
  const int b = blockIdx.x;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarps = nthr >> 5;
  const int Tb = min(seq_lens[b], T);
  const int L = label_lens[b];
  const int S = 2 * L + 1;
  const int blank = K - 1;
  const CtcSmem M = ctc_smem_layout(T, Lmax, K, ws.Sp);
  float* buf0 = smem + M.o_buf0;            // [Smax+2], entries 0,1 are NEG padding
  float* buf1 = smem + M.o_buf1;
  float* lse = smem + M.o_lse;              // [T] row log-sum-exp, log2 units
  int* lab = reinterpret_cast<int*>(smem + M.o_lab);      // [Lmax]  (reversed for DIR==1)
  float* red = smem + M.o_red;              // [32]
  double* coffs = reinterpret_cast<double*>(smem + M.o_coffs);   // cumulative renorm offset per block
  float* em = smem + M.o_em;                // [2][CTC_RB][kp] raw logits of the rows of a block
  float* stage = smem + M.o_stage;          // [2][CTC_RB][Sp] table rows of a block
  const int Kp = M.kp, Sp = ws.Sp;
  const bool staged = M.staged != 0;
  float* tab = (DIR == 0 ? ws.alpha : ws.beta) + (size_t)b * T * Sp;
  double* coff = ws.coff + ((size_t)b * 2 + DIR) * ws.NB;

  for (int i = tid; i < L; i += nthr) lab[i] = labels[(size_t)b * Lmax + (DIR == 0 ? i : L - 1 - i)];
  // row log-sum-exp: one THREAD per row so that every thread has a whole row of independent loads in flight
  for (int t = tid; t < Tb; t += nthr) {
    const float* row = logits + ((size_t)t * B + b) * K;
    float m = -INFINITY, s = 0.f;
    if (K <= 32) {
      float v[32];
#pragma unroll
      for (int k = 0; k < 32; ++k) v[k] = (k < K) ? __ldg(row + k) : -INFINITY;
#pragma unroll
      for (int k = 0; k < 32; ++k) m = fmaxf(m, v[k]);
#pragma unroll
      for (int k = 0; k < 32; ++k) s += __expf(v[k] - m);
    } else {
      for (int k = 0; k < K; ++k) m = fmaxf(m, __ldg(row + k));
      for (int k = 0; k < K; ++k) s += __expf(__ldg(row + k) - m);
    }
    lse[t] = (m + __logf(s)) * LOG2E;
  }
  if (tid < 2) { buf0[tid] = NEGF; buf1[tid] = NEGF; }
  // emissions of block 0
  {
    const int n0 = min(CTC_RB, Tb) * K;
    for (int e = tid; e < n0; e += nthr) {
      const int r = e / K, k = e - r * K;
      const int t = (DIR == 0) ? r : (Tb - 1 - r);
      em[r * Kp + k] = __ldg(logits + ((size_t)t * B + b) * K + k);
    }
  }
  __syncthreads();
  if (DIR == 0 && warp == 0) {
    // class-sorted (stable) permutation of label positions for the gradient kernel:
    // lane k owns class k (chunks of 32 classes), counting pass + warp scan + fill pass.
    int* cls = ws.cls_off + (size_t)b * (K + 1);
    int* perm = ws.perm + (size_t)b * Lmax;
    int base = 0;
    for (int k0 = 0; k0 < K; k0 += 32) {
      const int k = k0 + lane;
      int cnt = 0;
      for (int i = 0; i < L; ++i) cnt += (lab[i] == k);
      int incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      int c = base + incl - cnt;
      if (k < K) cls[k] = c;
      for (int i = 0; i < L; ++i)
        if (lab[i] == k) perm[c++] = i;
      base += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) cls[K] = base;
  }
  if (Tb <= 0) {
    if (DIR == 0 && tid == 0) { ws.logp[b] = 0.0; loss[b] = 0.f; }
    return;
  }

  // Per-thread lattice states s = tid + j*nthr: class, skip-transition flag and table column are loop-invariant.
  int cls_j[NS], sidx_j[NS];
  bool skip_j[NS], on_j[NS];
#pragma unroll
  for (int j = 0; j < NS; ++j) {
    const int s = tid + j * nthr;
    on_j[j] = s < S;
    cls_j[j] = (on_j[j] && (s & 1)) ? lab[s >> 1] : blank;
    skip_j[j] = on_j[j] && (s & 1) && s >= 2 && lab[s >> 1] != lab[(s >> 1) - 1];
    sidx_j[j] = (DIR == 0) ? s : (S - 1 - s);
  }
  const bool pre_in_regs = CTC_RB * K <= CTC_PRE * nthr;
  double cum = 0.0;                          // cumulative renorm offset (uniform across threads), log2 units
  float* prev = buf0;
  float* cur = buf1;
  if (tid == 0) coffs[0] = 0.0;
  const int nblk = (Tb + CTC_RB - 1) / CTC_RB;
  for (int blk = 0; blk < nblk; ++blk) {
    const int step0 = blk * CTC_RB;
    const int nsteps = min(CTC_RB, Tb - step0);
    const int nnext = min(CTC_RB, Tb - step0 - CTC_RB);      // rows of the next block (<= 0: none)
    const float* emc = em + (blk & 1) * CTC_RB * Kp;
    float* emn = em + ((blk + 1) & 1) * CTC_RB * Kp;
    float* stg = stage + (blk & 1) * CTC_RB * Sp;
    // next block's emissions -> registers (consumed at the end of this block)
    float pre[CTC_PRE];
    if (pre_in_regs) {
#pragma unroll
      for (int m = 0; m < CTC_PRE; ++m) {
        const int e = tid + m * nthr;
        pre[m] = 0.f;
        if (e < nnext * K) {
          const int r = e / K, k = e - r * K;
          const int st = step0 + CTC_RB + r;
          const int t = (DIR == 0) ? st : (Tb - 1 - st);
          pre[m] = __ldg(logits + ((size_t)t * B + b) * K + k);
        }
      }
    }
    for (int i = 0; i < nsteps; ++i) {
      const int step = step0 + i;
      const int t = (DIR == 0) ? step : (Tb - 1 - step);
      const float l = lse[t];
      const int row = (DIR == 0) ? i : (nsteps - 1 - i);     // staging rows ascend in t
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        if (on_j[j]) {
          const int s = tid + j * nthr;
          const float ly = fmaf(emc[i * Kp + cls_j[j]], LOG2E, -l);
          float v;
          if (step == 0) {
            v = (s < 2) ? 0.f : NEGF;
          } else {
            const float a2 = skip_j[j] ? prev[s] : NEGF;
            v = lse3(prev[s + 2], prev[s + 1], a2);
          }
          // alpha stores the value with the emission, beta stores it without; the running quantity carried to
          // the next step always includes the emission.  (NEGF = -1e30 absorbs every finite addend.)
          const float carried = v + ly;
          cur[s + 2] = carried;
          const float out = (DIR == 0) ? carried : v;
          if (staged) stg[row * Sp + sidx_j[j]] = out;
          else tab[(size_t)t * Sp + sidx_j[j]] = out;
        }
      }
      __syncthreads();
      float* tmp = prev; prev = cur; cur = tmp;
    }
    // ---- block boundary (prev holds the newest column)
    const bool last = (step0 + nsteps == Tb);
    if (!last) {
      float m = NEGF;
#pragma unroll
      for (int j = 0; j < NS; ++j)
        if (on_j[j]) m = fmaxf(m, prev[tid + j * nthr + 2]);
      m = warp_max(m);
      if (lane == 0) red[warp] = m;
      __syncthreads();
      float mm = red[0];
      for (int w = 1; w < nwarps; ++w) mm = fmaxf(mm, red[w]);
#pragma unroll
      for (int j = 0; j < NS; ++j)
        if (on_j[j]) prev[tid + j * nthr + 2] -= mm;
      cum += (double)mm;
      if (tid == 0) coffs[blk + 1] = cum;
      // park the next block's emissions
      if (pre_in_regs) {
#pragma unroll
        for (int m2 = 0; m2 < CTC_PRE; ++m2) {
          const int e = tid + m2 * nthr;
          if (e < nnext * K) {
            const int r = e / K, k = e - r * K;
            emn[r * Kp + k] = pre[m2];
          }
        }
      } else {
        for (int e = tid; e < nnext * K; e += nthr) {
          const int r = e / K, k = e - r * K;
          const int st = step0 + CTC_RB + r;
          const int t = (DIR == 0) ? st : (Tb - 1 - st);
          emn[r * Kp + k] = __ldg(logits + ((size_t)t * B + b) * K + k);
        }
      }
    } else if (DIR == 0 && tid == 0) {
      // log2 p = log2(2^alpha_{T-1}(S-1) + 2^alpha_{T-1}(S-2)) + offsets
      const float aL = prev[S - 1 + 2];
      const float aL1 = (S > 1) ? prev[S - 2 + 2] : NEGF;
      const float mx = fmaxf(aL, aL1);
      const double lp = (double)mx + (double)__log2f(exp2f(aL - mx) + exp2f(aL1 - mx)) + cum;
      ws.logp[b] = lp;
      loss[b] = (float)(-lp * LN2);
    }
    if (staged) fence_proxy_async();         // staged rows (generic proxy) -> visible to the bulk copy (async proxy)
    __syncthreads();
    if (staged) {
      if (tid == 0) {
        const int t_lo = (DIR == 0) ? step0 : (Tb - step0 - nsteps);
        bulk_store_s2g(tab + (size_t)t_lo * Sp, stg, (uint32_t)(nsteps * Sp * 4));
        bulk_commit();
        bulk_wait_read<1>();                 // the previous block's copy has finished reading its staging buffer
      }
      __syncthreads();
    }
  }
  if (tid == 0) {
    for (int j = 0; j < nblk; ++j) coff[j] = coffs[j];
    if (staged) bulk_wait_all();
  }
}

}  // namespace ds2
