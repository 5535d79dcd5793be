// Persistent clipped-ReLU recurrence, second exchange protocol (round 2), forward and backward.
//
// Same contraction layout as ds2_rnn_persistent.cu (src/custom_ops.py:48-72 inside tf.nn.bidirectional_dynamic_rnn,
// :91-93): units tiled by 128 (tcgen05 M), K split over a 4-CTA cluster, U slice resident in TMEM, 16-utterance
// instances as MMA N, partial sums reduced through DSMEM.  What changed is how a step's result reaches the other CTAs.
// Measured on B200 (tools/microbench/, profiles/r2_microbench_*.txt): with the round-1 protocol - rows stored into the
// [T,B,H] tensor, CTA barrier, one red on a counter, one poller, CTA barrier, 7 strong 16-byte loads per thread,
// st.shared, proxy fence, mbarrier hand-off to the MMA warp - the exchange alone costs ~3500 clk of a ~6100 clk step.
// Here:
//   * every CTA ALSO writes its 32 units x 16 utterances into an EXCHANGE RING in global memory that is laid out as
//     the shared-memory image the MMA wants (K-major, 128B swizzle): ring[direction][instance][slot][K rank] is the
//     complete B operand of one consumer, 14 KB contiguous;
//   * the thread that issues the MMAs polls the step counter itself, then fetches the image with ONE cp.async.bulk
//     straight into the operand buffer (no loader warps, no st.shared, no proxy fence, no hand-off) and issues the
//     MMAs as soon as the copy's mbarrier completes;
//   * publication has no CTA barrier: each epilogue warp adds 1 to the counter after its own stores;
//   * the step counter exists in several copies in different 4 KB blocks (see NCOPY below: L2 home die);
//   * the data still validates itself (there is no release fence: MEMBAR.GPU costs ~1000 clk a step): ring cells hold
//     the sentinel 0xFFFF until written, writers never emit that bf16 pattern, every cell is written by one 8-byte
//     strong store (st.relaxed.gpu.b64, single-copy atomic) and reset to the sentinel two steps before its slot is
//     reused; the epilogue warps check the landed image while the MMAs run and only take the accumulators when the
//     check has passed - a stale image is fetched again (never observed in practice; DS2_RNN_EXP=1 makes the check
//     fail once every 5th step so that the tests cover the re-fetch path);
//   * the [T,B,H] output is written but never read back by the kernel, so it needs no fill before the launch.
// DS2_RNN_V1=1 selects the round-1 kernel for A/B runs.
#include <stdlib.h>

#include "common.cuh"

namespace ds2 {

namespace p2 {

static constexpr int KS = 4;              // cluster size = K split
static constexpr int BP = 16;             // utterances per instance (MMA N)
static constexpr int INST_THREADS = 160;  // per instance: 1 control warp + 4 epilogue warps
static constexpr int MAX_NI = 4;
static constexpr int MAX_NBOX = 9;        // K slice <= 576 columns
static constexpr int RING = 8;            // exchange slots (>= 4: a slot is reset two steps before it is rewritten)
static constexpr int BOX_BYTES = BP * 128;   // one [16 x 64] bf16 box of the operand image
// Step counters exist in NCOPY copies, each in its own 4 KB block of the workspace.  B200's L2 is split over two dies and
// an address is homed on one of them (4 KB granularity, measured: tools/microbench/topo2.cu, profiles/r2_microbench_topo2.txt): a strong load or an atomic
// on a far-die address costs ~700 clk against ~300 clk on the near die, so a single counter makes half of the pollers
// pay the die crossing on every poll (counter barrier among 56 same-die CTAs: 1250 clk with a far counter, 650 clk with a near
// one).  Publishers add to every copy (fire and forget); each poller times the copies during its first steps and then
// polls the one that answers fastest.
static constexpr int NCOPY = 4;
static constexpr int CTR_COPY_STRIDE = 4096 / 4;     // in 32-bit words

struct Params {
  const float* add_in;     // fwd: wx [T,B,ND*H] ; bwd: dout [T,B,H]
  const float* bias;       // fwd: [ND,H]
  const int* lens;         // [B]
  const __nv_bfloat16* hs_in;   // bwd: hs [ND,T,B,H] (activation mask)
  float* dbias;            // bwd, optional: [ND,H]
  __nv_bfloat16* out;      // fwd: hs [ND,T,B,H] ; bwd: dz [T,B,ND*H]
  float* out32;            // fwd, optional: the same hidden states before their rounding to bf16, [ND,T,B,H] fp32
  unsigned char* ring;     // [ND][NI][RING][KS][nbox*BOX_BYTES]
  unsigned int* ctr;       // NCOPY x [ND][NI] step counters, 32 uints apart, copies CTR_COPY_STRIDE apart
  long long* prof;         // optional stage stamps of CTA (0,0,0) instance 0, 16 per step
  long out_row_pitch, out_dir_stride;
  int T, B, H, ND, reverse_mask;
  int NI, tiles_m, ksl, nbox;
  float scale, cap;
  int exp;                 // DS2_RNN_EXP bit 0: the image check reports 'stale' once on every 5th step (exercises the re-fetch path)
  int slab16;              // partial sums cross the cluster as bf16 (default) or fp32
  int poll_margin;         // clk before the predicted completion of a step at which its counter is first read (0: from the start)
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_async_v4(uint32_t remote_addr, uint32_t remote_bar, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%2, %3, %4, %5}, [%1];"
               ::"r"(remote_addr), "r"(remote_bar), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st_relaxed_b64(void* p, uint32_t lo, uint32_t hi) {
  asm volatile("{\n\t.reg .b64 v;\n\tmov.b64 v, {%1, %2};\n\tst.relaxed.gpu.global.b64 [%0], v;\n\t}" ::"l"(p), "r"(lo), "r"(hi) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t sdst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(sdst), "l"(gsrc), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_cp_128x256b(uint32_t taddr, uint64_t sdesc) {
  asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
// max over the 16-bit halves of two words (the sentinel 0xFFFF is the largest half-word: one VIMNMX per word instead of
// a compare per half)
__device__ __forceinline__ uint32_t max_u16x2(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
// clears the sign of a bf16 half that equals the sentinel (negative NaN, all payload bits set)
__device__ __forceinline__ uint32_t scrub_fill2(uint32_t w) { return w & ~(__vcmpeq2(w, 0xFFFFFFFFu) & 0x80008000u); }

// byte offset of element (utterance row n, slice column k) inside a consumer's operand image:
// box k/64 of [16 rows x 128 B], 16-byte chunk (k%64)/8 XOR-swizzled with the row (SWIZZLE_128B, K-major)
__host__ __device__ inline int image_offset(int n, int k) {
  return (k >> 6) * BOX_BYTES + n * 128 + ((((k & 63) >> 3) ^ (n & 7)) << 4) + (k & 7) * 2;
}

// shared-memory carve-up (offsets from a 1024-aligned base)
struct Smem {
  uint32_t a_off, b_off, slab_off, bar_off, flag_off, total;
};
static constexpr int BARS_PER_INST = 5;   // copy (tx), ok (4 warps), mma (commit), slab (tx), second half of the copy (tx)
__host__ __device__ inline Smem smem_layout(int nbox, int NI) {
  Smem s;
  s.a_off = 0;                                            // nbox x [128 x 64] bf16 staging of the U slice
  s.b_off = s.a_off + nbox * 16384;                       // NI x operand image (nbox boxes)
  s.slab_off = s.b_off + NI * nbox * BOX_BYTES;           // NI x KS x [32 x 16] fp32
  s.bar_off = s.slab_off + NI * KS * 32 * BP * 4;
  s.flag_off = s.bar_off + 8 * (2 + NI * BARS_PER_INST);  // tmem slot + per instance two "stale fetch" words
  s.total = s.flag_off + 16 + NI * 8;
  return s;
}

// all MMAs of one step: D[128 x 16] (+)= U_slice[128 x ksl] (TMEM) . h_slice^T (operand image in shared memory)
// boxes [B0, B1) of the image, unrolled (the image is fetched in two halves: the first half's MMAs issue while the second lands)
template <int B0, int B1>
__device__ __forceinline__ void issue_box_mmas(uint32_t d_tmem, uint32_t tmem_base, uint32_t sB_addr, uint32_t idesc) {
#pragma unroll
  for (int i = B0; i < B1; ++i) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      umma_bf16_ts(d_tmem, tmem_base + (i * 4 + k) * 8, umma_smem_desc_sw128(sB_addr + i * BOX_BYTES + k * 32, 16, 1024), idesc,
                   (i > 0 || k > 0) ? 1u : 0u);
  }
}
template <int NBOX>
__device__ __forceinline__ void issue_step_mmas(uint32_t d_tmem, uint32_t tmem_base, uint32_t sB_addr, uint32_t idesc, int nbox_rt) {
  if (NBOX > 0) {
#pragma unroll
    for (int i = 0; i < NBOX; ++i) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        umma_bf16_ts(d_tmem, tmem_base + (i * 4 + k) * 8, umma_smem_desc_sw128(sB_addr + i * BOX_BYTES + k * 32, 16, 1024), idesc,
                     (i > 0 || k > 0) ? 1u : 0u);
    }
  } else {
    for (int i = 0; i < nbox_rt; ++i) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        umma_bf16_ts(d_tmem, tmem_base + (i * 4 + k) * 8, umma_smem_desc_sw128(sB_addr + i * BOX_BYTES + k * 32, 16, 1024), idesc,
                     (i > 0 || k > 0) ? 1u : 0u);
    }
  }
}

// NIMAX bounds the block size (and so the registers per thread); NBOX > 0 unrolls the MMA issue for that slice length.
template <bool BWD, int NIMAX, int NBOX>
__global__ void __launch_bounds__(INST_THREADS * NIMAX, 1)
rnn2_kernel(const __grid_constant__ CUtensorMap tm_a, Params p) {
  extern __shared__ uint8_t rnn2_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(rnn2_smem_raw) + 1023) & ~uintptr_t(1023));
  const int NI = p.NI, nbox = NBOX > 0 ? NBOX : p.nbox;
  const Smem L = smem_layout(nbox, NI);
  uint8_t* sA = smem + L.a_off;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
  uint64_t* a_bar = bars;            // U slice landed in smem
  uint64_t* acp_bar = bars + 1;      // U slice copied into TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.flag_off);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int inst = warp / 5, w_in = warp % 5;
  // shared-space addresses used inside the step loop, computed once
  const uint32_t copy_bar_a = keep_reg(smem_u32(bars + 2 + inst * BARS_PER_INST));   // operand image landed (tx bytes)
  const uint32_t ok_bar_a = copy_bar_a + 8;                                          // the 4 epilogue warps have checked it
  const uint32_t mma_bar_a = copy_bar_a + 16;                                        // accumulators ready
  const uint32_t slab_bar_a = copy_bar_a + 24;                                       // partial slabs of this step arrived
  const uint32_t copy2_bar_a = copy_bar_a + 32;                                      // second half of the operand image landed
  const uint32_t flag_a = keep_reg(smem_u32(smem + L.flag_off + 16 + inst * 8));     // [id & 1] = id of a fetch found stale
  const uint32_t sB_addr = keep_reg(smem_u32(smem + L.b_off + inst * nbox * BOX_BYTES));
  const uint32_t slab_local = keep_reg(smem_u32(smem + L.slab_off + inst * KS * 32 * BP * 4));

  const uint32_t rank = cluster_ctarank();                  // K slice owned by this CTA, rows it finishes
  const int mt = blockIdx.y, d = blockIdx.z;
  const int m0 = mt * 128;
  const int rev = (p.reverse_mask >> d) & 1;
  const int T = p.T, H = p.H, ND = p.ND, Bn = p.B;
  unsigned int* ctr = p.ctr + (d * NI + inst) * 32;         // copy c at ctr + c * CTR_COPY_STRIDE
  const unsigned int arrivals_per_step = 4u * KS * (unsigned int)p.tiles_m;     // every epilogue warp of the direction
  const uint32_t img_bytes = (uint32_t)nbox * BOX_BYTES;
  const size_t slot_stride = (size_t)KS * img_bytes;
  unsigned char* ring_di = p.ring + (size_t)(d * NI + inst) * RING * slot_stride;
  const bool prof_cta = p.prof && (blockIdx.x + blockIdx.y + blockIdx.z == 0) && inst == 0;
  const int acols = nbox * 32;                              // TMEM columns holding U (8 per K=16 chunk)

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm_a);
    mbar_init(a_bar, 1);
    mbar_init(acp_bar, 1);
    for (int i = 0; i < NI; ++i) {
      uint64_t* bb = bars + 2 + i * BARS_PER_INST;
      mbar_init(bb, 1);            // arrive.expect_tx by the control thread + the bulk copy's bytes
      mbar_init(bb + 1, 4);        // one arrive per epilogue warp
      mbar_init(bb + 2, 1);        // tcgen05.commit
      mbar_init(bb + 3, 1);        // arrive.expect_tx per use + the st.async bytes
      mbar_init(bb + 4, 1);        // second half of the image
      uint32_t* fl = reinterpret_cast<uint32_t*>(smem + L.flag_off + 16 + i * 8);
      fl[0] = 0u;
      fl[1] = 0u;
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t d_tmem = tmem_base + acols + inst * BP;
  cluster_sync_all();      // every CTA of the cluster has initialised its barriers before any remote st.async

  if (w_in == 0) {
    // =========================== control warp of this instance: wait, fetch, MMA issue ===========================
    if (inst == 0 && elect_one_sync()) {
      mbar_arrive_expect_tx(a_bar, nbox * 16384);
      for (int i = 0; i < nbox; ++i) tma_load_2d(sA + i * 16384, &tm_a, a_bar, rank * p.ksl + i * 64, d * H + m0);
      mbar_wait(a_bar, 0, 10);
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA);
      for (int c = 0; c < nbox * 4; ++c)
        tmem_cp_128x256b(tmem_base + c * 8, umma_smem_desc_sw128(a0 + (c >> 2) * 16384 + (c & 3) * 32, 16, 1024));
      umma_commit(acp_bar);
    }
    __syncwarp();
    mbar_wait(acp_bar, 0, 11);       // U is resident in TMEM
    tc_fence_after();
    if (elect_one_sync()) {
      const uint32_t idesc = umma_idesc_bf16(128, BP, 0, 0);
      const unsigned char* my_img = ring_di + (size_t)rank * img_bytes;
      const unsigned char* src = my_img;
      uint32_t phc = 0;
      unsigned int round = 0;                    // fetches issued so far = id of the last one (0: none yet)
      const unsigned int* poll = ctr;            // the counter copy this thread polls
      long long best_rt = 0x7fffffffffffffffll;
      long long seen1 = 0, seen2 = 0;              // when the counter was seen complete in the last two steps
      long long pmin = 1ll << 40;                  // shortest step period so far
      long long wait_sum = 0, fetch_sum = 0;       // DS2_RNN_PROF: per-CTA totals (waiting for the others / fetch + MMA issue)
      // fetch the operand image at `src`, issue the step's MMAs, hand the accumulators over
      auto fetch_and_issue = [&]() {
        if (NBOX > 1) {
          // two halves on two barriers: the MMAs of the first boxes are issued while the rest of the image lands
          constexpr int H0 = NBOX > 1 ? (NBOX + 1) / 2 : 1, NB = NBOX > 1 ? NBOX : 2;
          mbar_arrive_expect_tx(copy_bar_a, H0 * BOX_BYTES);
          bulk_g2s(sB_addr, src, H0 * BOX_BYTES, copy_bar_a);
          mbar_arrive_expect_tx(copy2_bar_a, (NB - H0) * BOX_BYTES);
          bulk_g2s(sB_addr + H0 * BOX_BYTES, src + H0 * BOX_BYTES, (NB - H0) * BOX_BYTES, copy2_bar_a);
          mbar_wait(copy_bar_a, phc, 30);
          tc_fence_after();
          issue_box_mmas<0, H0>(d_tmem, tmem_base, sB_addr, idesc);
          mbar_wait(copy2_bar_a, phc, 30);
          tc_fence_after();
          issue_box_mmas<H0, NB>(d_tmem, tmem_base, sB_addr, idesc);
          phc ^= 1;
        } else {
          mbar_arrive_expect_tx(copy_bar_a, img_bytes);
          bulk_g2s(sB_addr, src, img_bytes, copy_bar_a);
          mbar_wait(copy_bar_a, phc, 30);
          phc ^= 1;
          tc_fence_after();
          issue_step_mmas<NBOX>(d_tmem, tmem_base, sB_addr, idesc, nbox);
        }
        umma_commit(mma_bar_a);
        ++round;
      };
      // step == T: nothing to fetch, but the last step's fetch may still have to be repeated
      for (int step = 1; step <= T; ++step) {
        const bool pr = prof_cta && step < 512;
        if (pr) p.prof[step * 16 + 0] = dev_clock();
        // every CTA of this direction has published step-1 of this instance
        const unsigned int need = arrivals_per_step * (unsigned int)step;
        const bool calib = step <= 2 * NCOPY;
        const unsigned int* q = calib ? ctr + ((step - 1) % NCOPY) * CTR_COPY_STRIDE : poll;
        {
          const long long t0 = dev_clock();
          int npoll = 1;
          SpinGuard g;
          if (p.poll_margin > 0 && step > 2 * NCOPY + 2) {
            // No polling while the step cannot be complete yet: 2 x 112 threads reading the counter lines back to back
            // slow down the increments that everybody is waiting for (measured: -4..5 % per step).  Polling starts
            // poll_margin clk before the moment the last period predicts - the estimate capped at 1.1 x the shortest
            // period seen, so that a late start (which lengthens that period) cannot feed back into later estimates.
            long long pe = seen1 - seen2;
            if (pe > pmin + pmin / 10) pe = pmin + pmin / 10;
            const long long start = seen1 + pe - (long long)p.poll_margin;
            while (dev_clock() < start) {}
          }
          while (ld_relaxed_gpu_u32(q) < need) {
            ++npoll;
            if (round > 0 && lds_u32(flag_a + 4 * (round & 1)) == round) {
              // The epilogue warps found a cell of the last image still holding the sentinel: the counter ran ahead of
              // the data.  Once all four have reported, fetch the same image again.
              mbar_wait(ok_bar_a, (round - 1) & 1, 31);
              __nanosleep(100);
              fetch_and_issue();
              if (pr) p.prof[step * 16 + 7] += 1;
            }
            g.check(20);
          }
          seen2 = seen1;
          seen1 = dev_clock();
          if (step > 2 * NCOPY + 1 && seen1 - seen2 < pmin) pmin = seen1 - seen2;   // a late start can lengthen a period, never shorten it
          if (calib) {
            const long long rt = (dev_clock() - t0) / npoll;
            if (rt < best_rt) { best_rt = rt; poll = q; }
          }
          if (p.prof && step > 2 * NCOPY) wait_sum += dev_clock() - t0;
        }
        if (step == T) break;
        if (pr) p.prof[step * 16 + 1] = dev_clock();
        const long long tf0 = p.prof ? dev_clock() : 0;
        src = my_img + (size_t)((step - 1) % RING) * slot_stride;
        fetch_and_issue();
        if (pr) p.prof[step * 16 + 4] = dev_clock();
        if (p.prof) fetch_sum += dev_clock() - tf0;
      }
      if (p.prof && inst == 0) {
        // per-CTA totals behind the per-step stamps of CTA 0: 4 values per CTA from slot 8192 on
        const int cta = (blockIdx.z * gridDim.y + blockIdx.y) * KS + (int)rank;
        unsigned int sm;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
        p.prof[8192 + cta * 4 + 0] = wait_sum;
        p.prof[8192 + cta * 4 + 1] = fetch_sum;
        p.prof[8192 + cta * 4 + 2] = (long long)sm;
        p.prof[8192 + cta * 4 + 3] = (long long)(poll - ctr) / CTR_COPY_STRIDE;
      }
    }
    __syncwarp();
  } else {
    // =========================== epilogue warps of this instance ===========================
    const int q = warp & 3;                      // TMEM lane quarter = rows [32q, 32q+32) of the tile
    const int et = (w_in - 1) * 32 + lane;       // 0..127
    const int bl = et % BP;                      // utterance within the instance
    const int jg = et / BP;                      // 0..7: 4-row group of this CTA's 32 rows
    const int b = inst * BP + bl;
    const int j = m0 + 32 * (int)rank + 4 * jg;  // first of 4 consecutive output units
    const bool jvalid = j < H;                   // H % 8 == 0 -> all-or-nothing
    float bias4[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) bias4[i] = (!BWD && jvalid) ? p.bias[d * H + j + i] : 0.f;
    const int len_b = p.lens[b];
    // where this thread's 4 units live in the exchange ring: the image of the K slice they belong to
    const int ks_p = jvalid ? j / p.ksl : 0;
    unsigned char* ring_cell = ring_di + (size_t)ks_p * img_bytes + image_offset(bl, jvalid ? j - ks_p * p.ksl : 0);
    const uint32_t slab_dst = map_to_rank(slab_local, (uint32_t)q) + ((rank * 32 + lane) * BP) * (p.slab16 ? 2 : 4);
    const uint32_t slab_rbar = map_to_rank(slab_bar_a, (uint32_t)q);
    const uint32_t slab_src = slab_local + ((4 * jg) * BP + bl) * (p.slab16 ? 2 : 4);
    const uint32_t chk_addr = sB_addr + et * 16;
    // global pointers that move by a constant per step
    const int t_first = BWD ? (rev ? 0 : T - 1) : (rev ? T - 1 : 0);
    const long t_inc = (BWD ? (rev ? 1 : -1) : (rev ? -1 : 1));
    const float* add_ptr = BWD ? p.add_in + ((size_t)t_first * Bn + b) * H + j
                               : p.add_in + ((size_t)t_first * Bn + b) * ND * H + (size_t)d * H + j;
    const long add_inc = t_inc * (long)Bn * (BWD ? H : ND * H);
    const __nv_bfloat16* hm_ptr = BWD ? p.hs_in + (((size_t)d * T + t_first) * Bn + b) * H + j : nullptr;
    const long hm_inc = t_inc * (long)Bn * H;
    __nv_bfloat16* out_ptr = p.out + (size_t)d * p.out_dir_stride + ((size_t)t_first * Bn + b) * p.out_row_pitch + j;
    const long out_inc = t_inc * (long)Bn * p.out_row_pitch;
    float* out32_ptr = (!BWD && p.out32) ? p.out32 + (((size_t)d * T + t_first) * Bn + b) * H + j : nullptr;
    float dbs[4] = {0.f, 0.f, 0.f, 0.f};         // backward: running sum over time of this thread's dz (bias gradient)
    uint32_t phc = 0, pho = 0, phm = 0, phs = 0, vround = 0;
    int t = t_first;
    for (int step = 0; step < T; ++step, t += (int)t_inc, add_ptr += add_inc, out_ptr += out_inc) {
      const bool pr = prof_cta && et == 0 && step < 512;
      // prefetch the additive term (s*Wx_t or dout_t) and, for the backward, h_t for the mask
      float4 add = make_float4(0.f, 0.f, 0.f, 0.f);
      uint2 hm = make_uint2(0u, 0u);
      if (jvalid) {
        add = __ldg(reinterpret_cast<const float4*>(add_ptr));
        if (BWD) {
          hm = *reinterpret_cast<const uint2*>(hm_ptr);
          hm_ptr += hm_inc;
        }
      }
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      if (step > 0) {
        if (et == 0) mbar_arrive_expect_tx(slab_bar_a, p.slab16 ? KS * 32 * BP * 2 : KS * 32 * BP * 4);   // arm this step's slab barrier
        // ---- check the operand image the control thread fetched (while the MMAs run), then take the accumulators
        bool force_stale = (p.exp & 1) && step % 5 == 0;
        for (;;) {
          mbar_wait(copy_bar_a, phc, 40);
          if (NBOX > 1) mbar_wait(copy2_bar_a, phc, 43);
          phc ^= 1;
          ++vround;
          if (pr) p.prof[step * 16 + 13] = dev_clock();
          uint32_t mx = 0u;
          if (NBOX > 0) {
#pragma unroll
            for (int m = 0; m < (NBOX > 0 ? NBOX : 1); ++m) {
              const uint4 v = lds_v4(chk_addr + m * 2048);
              mx = max_u16x2(mx, max_u16x2(max_u16x2(v.x, v.y), max_u16x2(v.z, v.w)));
            }
          } else {
            for (int m = 0; m < nbox; ++m) {
              const uint4 v = lds_v4(chk_addr + m * 2048);
              mx = max_u16x2(mx, max_u16x2(max_u16x2(v.x, v.y), max_u16x2(v.z, v.w)));
            }
          }
          const bool stale = ((mx & 0xFFFFu) == 0xFFFFu) || ((mx >> 16) == 0xFFFFu) || force_stale;
          force_stale = false;
          if (__any_sync(0xffffffffu, stale) && lane == 0) sts_u32(flag_a + 4 * (vround & 1), vround);
          if (lane == 0) mbar_arrive(ok_bar_a);
          if (pr) p.prof[step * 16 + 8] = dev_clock();
          mbar_wait(mma_bar_a, phm, 41);
          phm ^= 1;
          mbar_wait(ok_bar_a, pho, 42);          // all four warps have reported: the verdict is final
          pho ^= 1;
          if (lds_u32(flag_a + 4 * (vround & 1)) != vround) break;
        }
        tc_fence_after();
        if (pr) p.prof[step * 16 + 9] = dev_clock();
        // ---- partial sums of this CTA's K slice -> the CTA that finishes those rows
        uint32_t r[16];
        tmem_ld_32x16(d_tmem + (static_cast<uint32_t>(q * 32) << 16), r);
        tmem_ld_wait();
        tc_fence_before();
        if (p.slab16) {
          // the DSMEM store queue bounds this exchange: bf16 halves the stores (profiles/r1_rnn_stage_cycles_v6.txt)
          uint32_t h[8];
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(r[2 * c]), __uint_as_float(r[2 * c + 1]));
            h[c] = *reinterpret_cast<uint32_t*>(&t2);
          }
          st_async_v4(slab_dst, slab_rbar, h[0], h[1], h[2], h[3]);
          st_async_v4(slab_dst + 16, slab_rbar, h[4], h[5], h[6], h[7]);
        } else {
#pragma unroll
          for (int c = 0; c < 4; ++c) st_async_v4(slab_dst + c * 16, slab_rbar, r[4 * c], r[4 * c + 1], r[4 * c + 2], r[4 * c + 3]);
        }
        if (pr) p.prof[step * 16 + 10] = dev_clock();
        // ---- wait for the KS slabs of this step, reduce
        mbar_wait(slab_bar_a, phs, 50);
        phs ^= 1;
        if (pr) p.prof[step * 16 + 11] = dev_clock();
        if (p.slab16) {
#pragma unroll
          for (int s = 0; s < KS; ++s) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              unsigned short hv;
              asm volatile("ld.shared.u16 %0, [%1];" : "=h"(hv) : "r"(slab_src + (s * 32 + i) * BP * 2) : "memory");
              acc[i] += __uint_as_float((uint32_t)hv << 16);
            }
          }
        } else {
#pragma unroll
          for (int s = 0; s < KS; ++s) {
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[i] += lds_f32(slab_src + (s * 32 + i) * BP * 4);
          }
        }
      }
      float o[4] = {0.f, 0.f, 0.f, 0.f};
      uint2 pk = make_uint2(0u, 0u);
      if (jvalid) {
        const bool live = t < len_b;
        const float av[4] = {add.x, add.y, add.z, add.w};
        if (!BWD) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float z = fmaf(p.scale, av[i], acc[i] + bias4[i]);
            o[i] = live ? fminf(fmaxf(z, 0.f), p.cap) : 0.f;          // never NaN, never negative: no bf16 0xFFFF possible
          }
        } else {
          const __nv_bfloat16* hb = reinterpret_cast<const __nv_bfloat16*>(&hm);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float h = __bfloat162float(hb[i]);
            o[i] = (live && h > 0.f && h < p.cap) ? (av[i] + acc[i]) : 0.f;
            dbs[i] += o[i];
          }
        }
        __nv_bfloat162 b0v = __floats2bfloat162_rn(o[0], o[1]), b1v = __floats2bfloat162_rn(o[2], o[3]);
        pk.x = *reinterpret_cast<uint32_t*>(&b0v);
        pk.y = *reinterpret_cast<uint32_t*>(&b1v);
        if (BWD) {
          pk.x = scrub_fill2(pk.x);       // a NaN gradient must not look like the sentinel
          pk.y = scrub_fill2(pk.y);
        }
        // the exchange ring first: the other CTAs wait for it
        st_relaxed_b64(ring_cell + (size_t)(step % RING) * slot_stride, pk.x, pk.y);
      }
      // publish: one relaxed increment per counter copy after the warp's stores (no fence: the image validates itself)
      __syncwarp();
      if (lane < NCOPY) red_relaxed_gpu_add(ctr + lane * CTR_COPY_STRIDE, 1u);
      // afterwards what nobody inside the kernel waits for: the slot two steps ahead back to the sentinel, the caller's tensors
      if (jvalid) {
        st_relaxed_b64(ring_cell + (size_t)((step + 2) % RING) * slot_stride, 0xFFFFFFFFu, 0xFFFFFFFFu);
        *reinterpret_cast<uint2*>(out_ptr) = pk;
        if (!BWD && out32_ptr) {
          *reinterpret_cast<float4*>(out32_ptr) = make_float4(o[0], o[1], o[2], o[3]);
          out32_ptr += out_inc;            // fwd: the fp32 tensor has the pitch of the bf16 one
        }
      }
      if (pr) p.prof[step * 16 + 12] = dev_clock();
    }
    if (BWD && p.dbias != nullptr) {
      // bias gradient: sum over the instance's 16 utterances (16 consecutive lanes share a unit quad), one red per unit
#pragma unroll
      for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) dbs[i] += __shfl_xor_sync(0xffffffffu, dbs[i], o);
      }
      if (bl == 0 && jvalid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) atomicAdd(p.dbias + (size_t)d * H + j + i, dbs[i]);
      }
    }
  }
  // no CTA may exit while a peer can still write into its shared memory
  tc_fence_before();
  cluster_sync_all();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// Fills the exchange ring with the sentinel (cells that no CTA ever writes - slice columns beyond H - with zero, so
// that the MMA multiplies the zero-padded columns of U with zeros) and clears the step counters.
__global__ void __launch_bounds__(256) rnn2_ring_init_kernel(uint4* ring, long n_units, int nbox, int ksl, int H, unsigned int* ctr, int n_ctr) {
  const long tid = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (tid < n_ctr) ctr[tid] = 0u;
  const int units_per_img = nbox * (BOX_BYTES / 16);
  for (long u = tid; u < n_units; u += (long)gridDim.x * blockDim.x) {
    const long img = u / units_per_img;
    const int krank = (int)(img % KS);
    const int w = (int)(u - img * units_per_img);        // unit inside the image: box, row, chunk
    const int box = w >> 7, row = (w >> 3) & 15, chunk = w & 7;
    const int k = box * 64 + ((chunk ^ (row & 7)) << 3);
    const uint32_t v = (krank * ksl + k < H) ? 0xFFFFFFFFu : 0u;
    ring[u] = make_uint4(v, v, v, v);
  }
}

}  // namespace p2

// ------------------------------------------------------------------ host side
using namespace p2;

// implemented in ds2_rnn_persistent.cu (round-1 protocol, kept for A/B runs)
int rnn_relu_fwd_persist_v1(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs, int T, int B,
                            int H, int ND, int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st);
int rnn_relu_bwd_persist_v1(const float* dout, const void* u, const int* lens, const void* hs, void* dz, float* dbias, int T, int B, int H,
                            int ND, int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st);
size_t rnn_persist_workspace_bytes_v1(int T, int B, int H, int ND);

struct Geom {
  int tiles_m, ksl, nbox, ni;
  size_t smem, ring_bytes;
};

static int geometry(int B, int H, int ND, Geom* g) {
  if (B % BP != 0 || B < BP || B / BP > MAX_NI || H % 8 != 0 || H < 8 || (ND != 1 && ND != 2)) return 0;
  const int tm = (H + 127) / 128;
  const int slice = (H + KS - 1) / KS;
  const int l = ((slice + 63) / 64) * 64;
  const int nb = l / 64;
  if (nb > MAX_NBOX) return 0;
  const int n_inst = B / BP;
  if (nb * 32 + n_inst * BP > 512) return 0;   // TMEM columns: resident U + accumulators
  const size_t bytes = smem_layout(nb, n_inst).total + 1024;
  if (bytes > 225 * 1024) return 0;
  if (KS * tm * ND > 128) return 0;            // must be co-resident in clusters of 4 (148 SMs, 1 CTA / SM)
  if (g) {
    g->tiles_m = tm; g->ksl = l; g->nbox = nb; g->ni = n_inst; g->smem = bytes;
    g->ring_bytes = (size_t)ND * n_inst * RING * KS * nb * BOX_BYTES;
  }
  return 1;
}

int rnn_persist_supported(int B, int H, int ND) { return geometry(B, H, ND, nullptr); }

static constexpr size_t WS_CTR_BYTES = 2048;     // round-1 layout prefix: counters, stamps, U^T
static constexpr size_t WS_PROF_BYTES = 131072;

static size_t ring_offset(int H, int ND) {
  const size_t o = WS_CTR_BYTES + WS_PROF_BYTES + (size_t)ND * H * H * 2;
  return (o + 1023) & ~size_t(1023);
}

size_t rnn_persist_workspace_bytes(int T, int B, int H, int ND) {
  (void)T;
  Geom g;
  if (!geometry(B, H, ND, &g)) return rnn_persist_workspace_bytes_v1(T, B, H, ND);
  return ring_offset(H, ND) + g.ring_bytes + (size_t)(NCOPY + 1) * 4096;     // + the counter copies, each in its own 4 KB block
}

template <bool BWD, int NIMAX, int NBOX>
static int launch_nb(const CUtensorMap& ta, Params& p, size_t smem, cudaStream_t st) {
  auto kern = rnn2_kernel<BWD, NIMAX, NBOX>;
  static unsigned long long attr_done_devs = 0;
  if (first_use_on_device(&attr_done_devs)) {
    DS2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(KS, p.tiles_m, p.ND);
  cfg.blockDim = dim3(INST_THREADS * p.NI);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = KS;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;       // all CTAs co-resident (they wait on each other's data)
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = use_cooperative_launch() ? 2 : 1;    // profilers cannot replay cooperative launches
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, p);
  if (e != cudaSuccess && cfg.numAttrs == 2) {
    // some driver versions reject cooperative + cluster launches: the grid (<= 128 CTAs, 1 per SM) fits the device anyway
    (void)cudaGetLastError();
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, kern, ta, p);
  }
  if (e != cudaSuccess) return cuda_fail(e, BWD ? "rnn2_kernel<bwd>" : "rnn2_kernel<fwd>");
  count_launch();
  return DS2_OK;
}

template <bool BWD, int NIMAX>
static int launch_ni(const CUtensorMap& ta, Params& p, size_t smem, cudaStream_t st) {
  if (p.nbox == 7) return launch_nb<BWD, NIMAX, 7>(ta, p, smem, st);     // H = 1760 (train.sh)
  if (p.nbox == 4) return launch_nb<BWD, NIMAX, 4>(ta, p, smem, st);     // H = 800 .. 1024
  return launch_nb<BWD, NIMAX, 0>(ta, p, smem, st);
}

template <bool BWD>
static int launch(const CUtensorMap& ta, Params& p, size_t smem, cudaStream_t st) {
  return p.NI <= 2 ? launch_ni<BWD, 2>(ta, p, smem, st) : launch_ni<BWD, MAX_NI>(ta, p, smem, st);
}

static int common_setup(Params& p, Geom& g, int T, int B, int H, int ND, void* ws, size_t ws_bytes, const char* who, cudaStream_t st) {
  DS2_REQUIRE(geometry(B, H, ND, &g), DS2_ERR_UNSUPPORTED, "%s: persistent kernel does not take B=%d H=%d ND=%d", who, B, H, ND);
  const size_t need = rnn_persist_workspace_bytes(T, B, H, ND);
  DS2_REQUIRE(ws && ws_bytes >= need, DS2_ERR_WORKSPACE, "%s: workspace of %zu bytes required", who, need);
  DS2_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, DS2_ERR_BAD_ARG, "%s: workspace must be 256-byte aligned", who);
  p.tiles_m = g.tiles_m; p.ksl = g.ksl; p.nbox = g.nbox; p.NI = g.ni;
  p.ring = reinterpret_cast<unsigned char*>(ws) + ring_offset(H, ND);
  p.ctr = reinterpret_cast<unsigned int*>((reinterpret_cast<uintptr_t>(p.ring) + g.ring_bytes + 4095) & ~uintptr_t(4095));
  p.exp = env().rnn_exp;
  p.slab16 = env().rnn_slab_f32 ? 0 : 1;
  {
    // measured optimum per number of instances (B = 16, 32, 48, 64 at H = 1760): more chains, more jitter, earlier start;
    // too small a margin is worse than none (B = 64 at 650 clk: +10 %)
    static const int kMargin[MAX_NI + 1] = {0, 750, 725, 850, 900};
    p.poll_margin = env().rnn_poll_margin >= 0 ? env().rnn_poll_margin : kMargin[g.ni];
  }
  p.prof = env().rnn_prof ? reinterpret_cast<long long*>(static_cast<char*>(ws) + WS_CTR_BYTES) : nullptr;
  const long n_units = (long)(g.ring_bytes / 16);
  int blocks = (int)((n_units + 255) / 256);
  if (blocks > 4 * num_sms()) blocks = 4 * num_sms();
  rnn2_ring_init_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<uint4*>(p.ring), n_units, g.nbox, g.ksl, H, p.ctr, NCOPY * CTR_COPY_STRIDE);
  DS2_LAUNCH_CHECK("rnn2_ring_init_kernel");
  count_launch();
  return DS2_OK;
}

int rnn_relu_fwd_persist(const float* wx, float scale, const void* u, const float* bias, const int* lens, void* hs, float* hs32, int T, int B,
                         int H, int ND, int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (env().rnn_v1) {
    DS2_REQUIRE(!hs32, DS2_ERR_UNSUPPORTED, "ds2_rnn_relu_fwd_ex: the round-1 kernel (DS2_RNN_V1) has no fp32 output");
    return rnn_relu_fwd_persist_v1(wx, scale, u, bias, lens, hs, T, B, H, ND, reverse_mask, cap, ws, ws_bytes, st);
  }
  Params p = {};
  Geom g;
  int rc = common_setup(p, g, T, B, H, ND, ws, ws_bytes, "ds2_rnn_relu_fwd", st);
  if (rc) return rc;
  CUtensorMap ta;
  rc = make_tmap_2d_bf16(&ta, u, H, (uint64_t)ND * H, H, 64, 128);
  if (rc) return rc;
  p.add_in = wx; p.bias = bias; p.lens = lens; p.hs_in = nullptr; p.out = static_cast<__nv_bfloat16*>(hs);
  p.out32 = hs32;
  p.out_row_pitch = H; p.out_dir_stride = (long)T * B * H;
  p.T = T; p.B = B; p.H = H; p.ND = ND; p.reverse_mask = reverse_mask; p.scale = scale; p.cap = cap;
  return launch<false>(ta, p, g.smem, st);
}

// The backward contraction is U^T dz: `ut_in` (bf16 [ND,H,H], ut[d][k][j] = U[d][j][k]) is the resident A operand.  When
// the caller has no transposed copy (ut_in == nullptr) U is transposed into the workspace first.
int rnn_relu_bwd_persist(const float* dout, const void* u, const void* ut_in, const int* lens, const void* hs, void* dz, float* dbias, int T,
                         int B, int H, int ND, int reverse_mask, float cap, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (env().rnn_v1) {
    DS2_REQUIRE(u, DS2_ERR_BAD_ARG, "ds2_rnn_relu_bwd_ex: the round-1 kernel (DS2_RNN_V1) needs U itself");
    return rnn_relu_bwd_persist_v1(dout, u, lens, hs, dz, dbias, T, B, H, ND, reverse_mask, cap, ws, ws_bytes, st);
  }
  Params p = {};
  Geom g;
  int rc = common_setup(p, g, T, B, H, ND, ws, ws_bytes, "ds2_rnn_relu_bwd", st);
  if (rc) return rc;
  const void* ut = ut_in;
  if (!ut) {
    DS2_REQUIRE(u, DS2_ERR_BAD_ARG, "ds2_rnn_relu_bwd: neither U nor U^T given");
    __nv_bfloat16* utw = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(ws) + WS_CTR_BYTES + WS_PROF_BYTES);
    for (int d = 0; d < ND; ++d) {
      rc = ds2_transpose(static_cast<const __nv_bfloat16*>(u) + (size_t)d * H * H, DS2_BF16, H, H, H, utw + (size_t)d * H * H, DS2_BF16, H, st);
      if (rc) return rc;
    }
    ut = utw;
  }
  CUtensorMap ta;
  rc = make_tmap_2d_bf16(&ta, ut, H, (uint64_t)ND * H, H, 64, 128);
  if (rc) return rc;
  p.add_in = dout; p.bias = nullptr; p.lens = lens; p.hs_in = static_cast<const __nv_bfloat16*>(hs);
  p.dbias = dbias;
  if (dbias) DS2_CUDA(cudaMemsetAsync(dbias, 0, (size_t)ND * H * sizeof(float), st));
  p.out = static_cast<__nv_bfloat16*>(dz);
  p.out_row_pitch = (long)ND * H; p.out_dir_stride = H;
  p.T = T; p.B = B; p.H = H; p.ND = ND; p.reverse_mask = reverse_mask; p.scale = 1.f; p.cap = cap;
  return launch<true>(ta, p, g.smem, st);
}

}  // namespace ds2
