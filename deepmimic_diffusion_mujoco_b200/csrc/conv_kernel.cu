// K1 + K2: Conv1d-family implicit GEMM on tcgen05 (fp16 hi/lo split operands, fp32 accumulate in TMEM), TMA
// fed, with the reference block's epilogue fused: bias -> GroupNorm(8) -> Mish -> + time bias -> + residual.
// Two kernels share the device code below:
//   conv_gemm_kernel    one launch per layer (dm_unet_forward: arbitrary per-sample timesteps, layer debugging)
//   sample_loop_kernel  PERSISTENT: a CTA pair (tcgen05 cta_group::2) owns one 16-sample group and runs every layer of
//                       every denoising step plus the fused DDPM update (K4); samples never interact, so the whole
//                       p_sample_loop is one launch.  Template variants (the common one carries none of the others' code):
//                         RD  big tiles: TMEM rounds (H > 256) and wave-staged epilogues
//                         LS  L-split: clusters of four = two pairs per group, half of the positions each (small batches
//                             at H >= 128); GroupNorm sums and conv halos cross the split through distributed shared memory
//                         MC  weight multicast: clusters of four = two groups whose pairs share every weight tile
//                         AT  the attention blocks' CUDA-core stages (attn_stages.cuh) as jobs of the same list
//
// Data layout: activations are fp16 planes [plane][G][L][16][C] (plane 0 = hi, 1 = lo,
// value = hi + lo).  A CTA owns S = 8 samples of one group for all L positions (half of them when split), so its GEMM rows
// are ordered (l, b): row = l * S + b.  A conv tap that reads position l + d is then the SAME shared-memory
// tile viewed d * S rows further down -- a whole number of 8-row swizzle atoms -- so the five taps of a k=5
// conv are five UMMA descriptors over one TMA-loaded box (halo rows zero-filled by TMA out-of-bounds handling).
// The whole per-sample conv output lives in TMEM until GroupNorm statistics are known; no cross-CTA reduce.
//
// Issue discipline: the layer descriptor (ConvJob) is read from kernel-parameter space, the TMA and MMA warps
// run their loops warp-converged and issue through elect_one_sync(), so descriptors, coordinates and loop
// counters stay in uniform registers (no per-lane R2UR waterfall in front of every UTCHMMA / UTMALDG).  The
// 16 epilogue warps do not poll the accumulator barrier: one warp waits on the mbarrier, the others block on a
// hardware named barrier, which leaves the issue slots to the MMA / TMA warps.
//
// Reference: Conv1dBlock helpers.py:50-69; ResidualTemporalBlock temporal.py:37-46; Downsample1d/Upsample1d
// helpers.py:32-47; p_sample_loop diffusion.py:186-228 (under /root/reference/diffusion/diffuser/models/).
#include "attn_stages.cuh"
#include "conv_job.h"
#include "ptx.cuh"
#include "step_device.cuh"

#include <mutex>

namespace dm {

__device__ __forceinline__ float mish_f(float x) {
  // x * tanh(softplus(x)) == x * a / (a + 2), a = n (n + 2), n = e^x.  torch switches softplus to the identity
  // above 20; here the ratio already rounds to 1 there (a > 2^57), and the clamp keeps n finite.
  // Raw MUFU ops: __expf / __fdividef wrap ex2 / rcp in range fix-ups (3 extra instructions each) that cannot trigger
  // here: a + 2 >= 2, and an e^x that underflows (x < -87) gives mish = 0 where the exact value is below 1e-36.
  float n, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(n) : "f"(fminf(x, 30.f) * 1.4426950408889634f));
  const float a = n * (n + 2.f);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a + 2.f));
  return x * (a * r);
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kMaxDynSmem = 201 * 1024;   // dynamic shared memory opt-in of both kernels (plan.cu budgets 200 KB + alignment slack)
constexpr int kPollWarp = 2;   // epilogue warp that waits on the accumulator mbarrier (SMSP 2: neither TMA nor MMA warp)

// Optional phase trace (debug): tr points at kTraceSlots slots of this (job, unit); null in production.
// Slots 0-15: phases; 16 + 2i / 17 + 2i: issue of weight load i by the producer / its arrival seen by the MMA warp.
constexpr int kTraceSlots = 64;
// TR = false compiles every mark away: the production kernel carries no trace code (it costs ~3 % of the loop
// time through instruction-cache footprint alone); dm_plan_set_trace selects the TR = true instantiation.
template <bool TR>
__device__ __forceinline__ void trace_mark(unsigned long long* tr, int ev) {
  if (TR && tr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    tr[ev] = t;
  }
}

struct ConvShared {
  uint64_t a_full[2], a_empty[2], w_full[8], w_empty[8], acc_full, res_full;
  uint64_t tile_ready;   // resident link (leader CTA's copy is the live one): both CTAs' staged tiles are complete
  uint64_t halo_free;    // L-split resident links: the TWIN pair's MMAs of the current unit have completed (its halo rows may be overwritten)
  uint64_t mc_done;      // weight multicast: the MMAs of the current unit have completed in BOTH pairs of the cluster
  uint64_t xstat_full;   // L-split: the twin CTA (other half of L, same samples) has delivered its GroupNorm partial sums
  uint32_t tmem_slot;
  alignas(8) float xstat[2][128][2];    // L-split: the twin's (sum, sum of squares) per (sample, group), double-buffered by use
  alignas(16) float par[5][256];        // bias, gamma, beta, time bias, auxiliary-conv bias of this unit's Nc channels
  alignas(16) unsigned char qgrp[64];   // GroupNorm group (within the unit) of every 4-channel quad
  float stat[16][8][2];                 // per (sample, group): mean, rstd
  // per-warp partial sums [warp][sample][group][2], combined in a fixed order (deterministic); after the
  // combine the same storage holds the per-(sample, channel) GroupNorm scale/shift table (kTabStride floats/row)
  alignas(16) float part[kEpiWarps * 256 + 64];   // table row = 2 * Nc floats + 4 floats of bank skew, Nc <= 128
};

// Parity bookkeeping of the mbarrier rings.  Bit s of a mask = parity of the number of times stage s has been
// used; producer and MMA issuer walk the same stage sequence, so their private copies stay in step even
// though ring depths change from layer to layer inside the persistent kernel.
struct RoleState {
  uint32_t a_bits = 0, w_bits = 0, acc_count = 0, res_count = 0, tile_count = 0, x_count = 0;
  uint32_t mc_count = 0, h_count = 0;
  int pre_w = 0;   // producer: weight loads of the current unit that were already issued during the previous epilogue
  int lhalf = 0;       // L-split: which half of the positions this CTA's pair works on (0 when the job is not split)
  int pair_base = 0;   // rank of the pair's leader inside the cluster (0, or 2 for the second pair of a cluster of four)
};

// mc: weight stages are filled for, and released by, both pairs of the cluster (two MMA warps each)
__device__ __forceinline__ void init_shared(ConvShared& sh, bool mc = false, bool ls = false) {
  for (int i = 0; i < 2; ++i) { mbar_init(&sh.a_full[i], 1); mbar_init(&sh.a_empty[i], 2); }   // both MMA warps release
  for (int i = 0; i < 8; ++i) { mbar_init(&sh.w_full[i], 1); mbar_init(&sh.w_empty[i], mc ? 4 : 2); }
  mbar_init(&sh.mc_done, 4);
  mbar_init(&sh.acc_full, 2);
  mbar_init(&sh.res_full, 1);
  mbar_init(&sh.tile_ready, ls ? 4 : 2);   // one arrival from each CTA of the pair (the peer's is remote); L-split: and of the twin pair
  mbar_init(&sh.halo_free, 2);             // both issuer warps of the twin pair's leader
  mbar_init(&sh.xstat_full, 128);  // one remote arrival per (sample, group) value pair the twin writes
  fence_barrier_init();
}

// ================================ TMA producer (one warp per CTA, one elected lane issues) ==================
// Both CTAs of the pair run this loop: each loads ITS samples' activation rows and ITS half of every weight tile into
// its own shared memory.  The bytes of both CTAs are counted on the LEADER's full barriers (the leader's MMA warps are
// the only waiters): the leader arms them with arrive.expect_tx for twice its own byte count, the peer only issues its
// loads.  Stage reuse is gated per CTA by its own empty barriers, which the MMA commits signal in both CTAs.
// MC (weight multicast, clusters of four = two pairs working on two groups): the same-ranked CTAs of the two pairs need
// the same half of every weight tile, so each loads a QUARTER (two boxes, one per plane) and multicasts it to both.  A
// stage is refilled when BOTH pairs have released it (w_empty counts four), and this warp leaves a unit only when both
// pairs' MMAs have completed (mc_done): the next unit's fills land in the other pair's shared memory too.
__device__ __forceinline__ void load_w_quarter(const ConvJob& jb, const CUtensorMap* wq, uint32_t stage, uint64_t* bar, int wn0,
                                               int chunk, int tap, int rank, int pair) {
  const int rows_q = jb.Nc >> 2;
  const unsigned short mask = (unsigned short)((1u << rank) | (4u << rank));
  const uint32_t dst = stage + (uint32_t)(pair * rows_q * jb.row_bytes);
  tma_load_5d_pair_mc(dst, wq, bar, 0, wn0 + pair * rows_q, chunk, tap, 0, mask);
  tma_load_5d_pair_mc(dst + (uint32_t)jb.w_plane_bytes, wq, bar, 0, wn0 + pair * rows_q, chunk, tap, 1, mask);
}

template <bool TR, bool LS, bool MC>
__device__ __forceinline__ void producer_unit(const ConvJob& job, const ConvMaps* maps, const CUtensorMap* a_maps, const CUtensorMap* w_map,
                                              const CUtensorMap* x_map, int g, int b0, int n0, int rank, uint32_t a_smem,
                                              uint32_t w_smem, ConvShared& sh, RoleState& rs, unsigned long long* tr,
                                              uint32_t smem_base, const ConvJob* nj, const CUtensorMap* n_w_map, int n_n0,
                                              const CUtensorMap* n_wq_map) {
  const bool leader = elect_one_sync();
  const bool arm = leader && rank == 0;
  const int wn0 = n0 + rank * (job.Nc >> 1);          // this CTA's rows of the weight tile
  int as = 0, ws = 0, wi = 0;
  if (leader) trace_mark<TR>(tr, 2);
  for (int lt = 0; lt < job.n_ltiles; ++lt) {
    const int lstart = job.l_stride * (lt * job.Lt) - job.halo_lo + (LS ? rs.lhalf * job.ls_in : 0);
    int chunk0 = 0;   // index of the source's first channel chunk inside the concatenated input (weights are chunk-major)
    for (int src = 0; src < job.nsrc; ++src) {
      const int nch = job.src_chunks[src];
      for (int ch = 0; ch < nch; ++ch) {
        if (!job.a_res) {
          mbar_wait(&sh.a_empty[as], ((rs.a_bits >> as) & 1) ^ 1, 0x100 + as);
          if (arm) mbar_arrive_expect_tx(&sh.a_full[as], 2u * job.a_tx_bytes);
          if (leader)
            tma_load_5d_pair(a_smem + as * job.a_stage_bytes, &a_maps[src], &sh.a_full[as], ch * job.Cc, b0, lstart, g, 0);
          rs.a_bits ^= 1u << as;
        }
        for (int t = 0; t < job.ntaps + job.aux; ++t) {   // the auxiliary 1x1 weights ride the same ring as one more tap
          if (rs.pre_w > 0) {
            --rs.pre_w;            // issued during the previous layer's epilogue
          } else {
            mbar_wait(&sh.w_empty[ws], ((rs.w_bits >> ws) & 1) ^ 1, 0x200 + ws);
            if (arm) mbar_arrive_expect_tx(&sh.w_full[ws], 2u * job.w_tx_bytes);
            if (MC && leader) {
              load_w_quarter(job, t < job.ntaps ? &maps->wq_map : &maps->xq_map, w_smem + ws * job.w_stage_bytes, &sh.w_full[ws],
                             wn0, chunk0 + ch, t < job.ntaps ? job.taps[t].wtap : 0, rank, rs.pair_base >> 1);
            } else if (leader) {
              tma_load_5d_pair(w_smem + ws * job.w_stage_bytes, t < job.ntaps ? w_map : x_map, &sh.w_full[ws], 0, wn0,
                               chunk0 + ch, t < job.ntaps ? job.taps[t].wtap : 0, 0);
#if defined(DM_TRACE_WLOADS) && !defined(DM_TRACE_MMA)
              if (tr && wi < 24) trace_mark<TR>(tr, 16 + 2 * wi);
#endif
            }
            rs.w_bits ^= 1u << ws;
          }
          if (++ws == job.n_w_stages) ws = 0;
          ++wi;
        }
        if (!job.a_res && ++as == job.n_a_stages) as = 0;
      }
      chunk0 += nch;
    }
    // TMEM rounds: this warp observes EVERY accumulator phase in order (a parity wait is only valid one phase behind the
    // barrier), so it does not run ahead of a round boundary -- the MMA warps stop there anyway until the epilogues have
    // drained the round
    if (job.n_rounds > 1 && lt + 1 < job.n_ltiles && (lt + 1) % job.lt_round == 0) {
      mbar_wait(&sh.acc_full, rs.acc_count & 1, 0x720);
      rs.acc_count++;
    }
  }
  if (leader) trace_mark<TR>(tr, 4);
  if (MC) {   // both pairs' MMAs of this unit are complete: every operand stage of the unit is dead in all four CTAs
    mbar_wait(&sh.mc_done, rs.mc_count & 1, 0x730);
    rs.mc_count++;
  }
  // A resident link has no block barrier behind it (the roles run ahead, see sample_loop_kernel): this warp may only
  // touch the next layer's ring once this layer's MMAs have completed, whether or not it prefetches.
  if (job.next_res && !(nj != nullptr && job.pre_next > 0)) mbar_wait(&sh.acc_full, rs.acc_count & 1, 0x710);
  // Weight prefetch across the layer boundary: weights do not depend on activations.  Once this unit's MMAs
  // have all completed (its operand stages are dead), the first weight tiles of the next layer are loaded into
  // that layer's ring slots that lie beyond this layer's staged output tile; they land during the epilogue.
  if (nj != nullptr && job.pre_next > 0) {
    mbar_wait(&sh.acc_full, rs.acc_count & 1, 0x700);   // (the last round's phase; earlier ones were observed above)
    const uint32_t nw_smem = smem_base + (uint32_t)(nj->a_res ? nj->w_off_res : nj->n_a_stages * nj->a_stage_bytes);
    for (int i = 0; i < job.pre_next; ++i) {
      mbar_wait(&sh.w_empty[i], ((rs.w_bits >> i) & 1) ^ 1, 0x780 + i);
      if (arm) mbar_arrive_expect_tx(&sh.w_full[i], 2u * nj->w_tx_bytes);
      if (MC && leader)
        load_w_quarter(*nj, n_wq_map, nw_smem + i * nj->w_stage_bytes, &sh.w_full[i], n_n0 + rank * (nj->Nc >> 1), 0,
                       nj->taps[i].wtap, rank, rs.pair_base >> 1);
      else if (leader)
        tma_load_5d_pair(nw_smem + i * nj->w_stage_bytes, n_w_map, &sh.w_full[i], 0, n_n0 + rank * (nj->Nc >> 1), 0,
                         nj->taps[i].wtap, 0);
      rs.w_bits ^= 1u << i;
    }
    rs.pre_w = job.pre_next;
  }
  rs.acc_count++;
  __syncwarp();
}

// ================================ MMA issuers (leader CTA only: two warps, one elected lane each) =============
// The MMAs are CTA-pair instructions (tcgen05.mma.cta_group::2): one instruction multiplies both CTAs' activation rows
// (M = 2 * mrows; each CTA's rows come from its own shared memory and accumulate in its own TMEM) by a weight tile of
// which each CTA staged half the N rows.  Per SM an MMA then reads its A rows plus HALF of B from shared memory -- the
// single-CTA form reads all of B and is bound by the 128 B/clk shared-memory port, which the TMA writes also use.
//
// A single thread cannot keep the tensor pipe busy: every mbarrier poll or descriptor recomputation between two
// taps shows up as a pipe bubble (measured: 86-110 cycles per N=128 MMA from one thread against the 64-cycle MAC
// floor; two threads reach 64.4).  So two warps of the leader issue.  job.mma_mode:
//   0  warp A issues everything (B only keeps the ring bookkeeping in step);
//   1  every tap is shared: A issues the first half of its k-steps, B the second half, each into its OWN
//      accumulator set (B's at column job.acc2_off); the epilogue adds the two.  Every accumulator sees one
//      thread's MMAs in program order, so results stay bit-reproducible.  The auxiliary tap (one accumulator
//      set) is B's alone.
// A weight stage is released (w_empty, two arrivals, in BOTH CTAs) by both warps; A stages and the unit's acc_full
// take one commit from each warp as well.
// RD = false: the kernel variant for plans whose every layer fits TMEM at once (all horizons <= 256): the round logic
// compiles away.  RD = true: layers may pass through TMEM in rounds (job.n_rounds > 1).
template <bool TR, bool RD, bool LS, bool MC>
__device__ __forceinline__ void mma_unit(const ConvJob& job, uint32_t smem_base, uint32_t a_smem, uint32_t w_smem,
                                         ConvShared& sh, uint32_t tmem_base, RoleState& rs, unsigned long long* tr,
                                         const int which) {
  const bool leader = elect_one_sync();
  if (which != 0) tr = nullptr;
  if (leader) trace_mark<TR>(tr, 20);
  const int S = job.S, Nc = job.Nc, ncols = job.ncols;
  const int MT = (RD ? job.lt_round : job.n_ltiles) * job.Mt;      // tiles per phase resident in TMEM (one round's worth)
  const int mode = job.mma_mode;
  const uint32_t idesc = umma_idesc_f16(2 * job.mrows, Nc);
  const int ksteps = job.ksteps;   // 16-channel k-steps per chunk that hold real input channels (zero padding skipped)
  const int pos_per_mt = job.mrows / S;
  // descriptor high words are per-layer constants; only the 14-bit start address changes per MMA
  const uint64_t hi_a = umma_smem_desc(0, job.l_stride * 8 * job.row_bytes, job.swz);
  const uint64_t hi_w = umma_smem_desc(0, 8 * job.row_bytes, job.swz);
  const uint32_t a_plane16 = (uint32_t)(job.a_res ? job.a_res_plane : job.a_plane_bytes) >> 4;
  const uint32_t w_plane16 = job.w_plane_bytes >> 4;
  const uint32_t mt_step16 = (uint32_t)(job.l_stride * pos_per_mt * S * job.row_bytes) >> 4;
  const int pair_base = (LS || MC) ? rs.pair_base : 0;   // (a compile-time 0 in the common kernels)
  const uint32_t peer_w_empty = mapa_u32(smem_u32(&sh.w_empty[0]), (uint32_t)pair_base + 1);   // the peer CTA's copy of the ring's empty barriers
  const unsigned short cmask = (unsigned short)(3u << pair_base);     // the pair's ranks, for the multicast commits
  const unsigned short wmask = MC ? (unsigned short)0xF : cmask;      // weight stages are shared with the other pair
  const uint32_t twin_w_empty0 = MC ? mapa_u32(smem_u32(&sh.w_empty[0]), (uint32_t)(pair_base ^ 2)) : 0u;
  const uint32_t twin_w_empty1 = MC ? mapa_u32(smem_u32(&sh.w_empty[0]), (uint32_t)(pair_base ^ 2) + 1) : 0u;
  int as = 0, ws = 0, wi = 0;
  if (leader) trace_mark<TR>(tr, 21);
  int lt_in = 0;            // l-tile index inside the current TMEM round
  for (int lt = 0; lt < job.n_ltiles; ++lt) {
    int cidx = 0;
    uint32_t started = 0;   // mode 2: accumulator sets (phase 0, phase 1, auxiliary) this warp has written in this l-tile
    for (int src = 0; src < job.nsrc; ++src) {
      const int nch = job.src_chunks[src];
      for (int ch = 0; ch < nch; ++ch, ++cidx) {
        if (!job.a_res) {
          mbar_wait(&sh.a_full[as], (rs.a_bits >> as) & 1, 0x300 + as);
          rs.a_bits ^= 1u << as;
          tc_fence_after();
        } else if (lt == 0 && cidx == 0) {
          // resident link: this warp ran ahead of the previous layer's epilogue (descriptors and loop state are set up,
          // the first weight tiles were prefetched); the tiles it reads in place -- and the TMEM columns it overwrites,
          // in both CTAs -- are handed over by one mbarrier arrival per CTA instead of a barrier
          mbar_wait_cluster(&sh.tile_ready, rs.tile_count & 1, 0x310);
          rs.tile_count++;
          tc_fence_after();
        }
        if (leader && lt == 0 && cidx == 0) trace_mark<TR>(tr, 5);
#ifdef DM_TRACE_WLOADS
        if (leader && tr && lt == 0 && cidx < 8) trace_mark<TR>(tr, 48 + cidx);
#endif
        const uint32_t a_base16 =
            (job.a_res ? smem_base + (uint32_t)(job.a_res_off + cidx * job.a_res_chunk) : a_smem + as * job.a_stage_bytes) >> 4;
        for (int t = 0; t < job.ntaps + job.aux; ++t) {
          const bool is_aux = t >= job.ntaps;
          // mode 1: both warps work on every tap, A on the first half of its k-steps and B on the second half;
          // the auxiliary tap (single accumulator set) is B's alone.  mode 2: whole taps alternate between the warps
          // (a warp's per-tap bookkeeping -- commit, ring step, barrier poll, descriptors -- then overlaps the other
          // warp's issue: the layers with little MMA work per tap are bound by that bookkeeping).  mode 0: all A's.
          const bool both = mode == 1 && !is_aux;
          const bool mine = mode == 2 ? (wi & 1) == which : (both || which == mode);
          const uint32_t w_par = (rs.w_bits >> ws) & 1;
          rs.w_bits ^= 1u << ws;
          if (mode == 2) {
            // Whole taps alternate.  The ring depth is even, so a stage always serves the same warp: the other warp
            // neither waits for its fills nor releases it (the owner releases it twice); it only keeps its parity
            // bookkeeping in step.  Its next wait on that stage happens in a later layer, behind a hand-over that
            // implies every fill of this layer has completed (a parity wait may only be one phase behind the barrier).
            if (!mine) { ++wi; if (++ws == job.n_w_stages) ws = 0; continue; }
            mbar_wait(&sh.w_full[ws], w_par, 0x400 + ws);
          } else {
            // Both warps observe every fill of every stage, in order, and the stage is not refilled before both have:
            // a warp without work on a tap still waits for its fill and then releases its half of w_empty
            // (otherwise its half is the tcgen05.commit).
            mbar_wait(&sh.w_full[ws], w_par, 0x400 + ws);
          }
          if (!mine) {
            if (leader) {
              mbar_arrive(&sh.w_empty[ws]);
              mbar_arrive_remote(peer_w_empty + 8u * (uint32_t)ws);
              if (MC) {
                mbar_arrive_remote(twin_w_empty0 + 8u * (uint32_t)ws);
                mbar_arrive_remote(twin_w_empty1 + 8u * (uint32_t)ws);
              }
            }
          } else {
            tc_fence_after();
            if (leader && lt == 0 && cidx == 0 && t == 0) trace_mark<TR>(tr, 6);
#ifdef DM_TRACE_MMA
            if (leader && tr && wi < 24) tr[17 + 2 * wi] = (unsigned long long)clock64();
#elif defined(DM_TRACE_WLOADS)
            if (leader && tr && wi < 24) trace_mark<TR>(tr, 17 + 2 * wi);
#endif
            const uint32_t w16 = (w_smem + ws * job.w_stage_bytes) >> 4;
            const int shift = is_aux ? job.aux_shift : job.taps[t].shift, phase = is_aux ? 0 : job.taps[t].phase;
            bool first = cidx == 0 && (is_aux || job.taps[t].first != 0);
            if (mode == 2) {          // this warp's first product into (its accumulator set of) this phase clears it
              const uint32_t bit = is_aux ? 4u : 1u << phase;
              first = (started & bit) == 0;
              started |= bit;
            }
            const int k_mid = (ksteps + 1) >> 1;
            const int k0 = both && which ? k_mid : 0, k1 = both && !which ? k_mid : ksteps;
            uint32_t d = tmem_base + (uint32_t)((phase * MT + (RD ? lt_in : lt) * job.Mt) * ncols) +
                         (is_aux ? (uint32_t)job.aux_col0 : which ? (uint32_t)job.acc2_off : 0u);
            uint32_t a16 = a_base16 + ((uint32_t)(shift * S * job.row_bytes) >> 4);
            for (int mt = 0; mt < job.Mt; ++mt, d += ncols, a16 += mt_step16) {
              if (leader) {
                for (int k = k0; k < k1; ++k) {
                  const uint64_t a_hi = hi_a | ((a16 + 2 * k) & 0x3FFF);
                  const uint64_t a_lo = hi_a | ((a16 + a_plane16 + 2 * k) & 0x3FFF);
                  const uint64_t w_hi = hi_w | ((w16 + 2 * k) & 0x3FFF);
                  const uint64_t w_lo = hi_w | ((w16 + w_plane16 + 2 * k) & 0x3FFF);
                  umma_f16_pair(d, a_lo, w_hi, idesc, !(first && k == k0));
                  umma_f16_pair(d, a_hi, w_lo, idesc, 1);
                  umma_f16_pair(d, a_hi, w_hi, idesc, 1);
                }
              }
            }
#ifdef DM_TRACE_MMA   // debug: when this warp finished ISSUING the tap's MMAs (slot pairs: issue done / fill seen)
            if (leader && tr && wi < 24) tr[16 + 2 * wi] = (unsigned long long)clock64();
#endif
            __syncwarp();
            if (leader) {
              umma_commit_pair(&sh.w_empty[ws], wmask);
              if (mode == 2) umma_commit_pair(&sh.w_empty[ws], wmask);   // the owner's release counts for both warps
            }
          }
          ++wi;
          if (++ws == job.n_w_stages) ws = 0;
        }
        if (!job.a_res) {
          if (leader) umma_commit_pair(&sh.a_empty[as], cmask);
          if (++as == job.n_a_stages) as = 0;
        }
      }
    }
    if (!RD) {
      if (lt + 1 == job.n_ltiles && leader) {
        umma_commit_pair(&sh.acc_full, cmask);
        if (MC) umma_commit_pair(&sh.mc_done, (unsigned short)0xF);   // (one from each issuer warp of each pair)
        if (LS) umma_commit_pair(&sh.halo_free, (unsigned short)(3u << (pair_base ^ 2)));   // the twin pair may overwrite our halo rows
      }
    } else if (++lt_in == job.lt_round || lt + 1 == job.n_ltiles) {   // end of a TMEM round: hand the accumulators to the epilogues
      lt_in = 0;
      if (leader) umma_commit_pair(&sh.acc_full, cmask);
      if (MC && leader && lt + 1 == job.n_ltiles) umma_commit_pair(&sh.mc_done, (unsigned short)0xF);
      if (LS && leader && lt + 1 == job.n_ltiles) umma_commit_pair(&sh.halo_free, (unsigned short)(3u << (pair_base ^ 2)));
      if (lt + 1 < job.n_ltiles) {
        // the next round overwrites these columns in BOTH CTAs: wait until both epilogues have drained them
        mbar_wait_cluster(&sh.tile_ready, rs.tile_count & 1, 0x320);
        rs.tile_count++;
        tc_fence_after();
      }
    }
  }
  if (leader) trace_mark<TR>(tr, 7);
  __syncwarp();
}

// Cross-lane part of the GroupNorm statistics.  Each lane holds NG (sum, sum of squares) pairs, one per
// GroupNorm group of its 32 channels; lanes with the same sample index (lane % S) must be added.  Recursive
// halving: at every butterfly step a lane keeps half of its values and hands the other half to its partner, so
// NG values cost NG + NG/2 shuffles per array instead of 2 * NG, and the final smem accumulation is spread over
// all lanes.  Every (warp, sample, group) slot is touched by exactly one lane per call: the order is fixed.
template <int NG, bool S8>
__device__ __forceinline__ void gn_reduce_rows(float* s, float* q, int lane, float* part_warp, const unsigned char* qg) {
  constexpr int QS = 8 / NG;   // 4-channel quads per folded value
  auto rmw = [&](int bsample, int r, float sv, float qv) {
    float2* pp = reinterpret_cast<float2*>(&part_warp[(bsample * 8 + qg[r * QS]) * 2]);
    float2 acc = *pp;
    acc.x += sv;
    acc.y += qv;
    *pp = acc;
  };
  if constexpr (NG >= 2) {
    constexpr int H = NG / 2;
    const bool h1 = (lane & 16) != 0;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const float ss = h1 ? s[i] : s[i + H], sk = h1 ? s[i + H] : s[i];
      const float qs = h1 ? q[i] : q[i + H], qk = h1 ? q[i + H] : q[i];
      s[i] = sk + __shfl_xor_sync(0xffffffffu, ss, 16);
      q[i] = qk + __shfl_xor_sync(0xffffffffu, qs, 16);
    }
    const int r1 = h1 ? H : 0;
    if constexpr (S8) {
      const bool h0 = (lane & 8) != 0;
      if constexpr (H >= 2) {
        constexpr int H2 = H / 2;
#pragma unroll
        for (int i = 0; i < H2; ++i) {
          const float ss = h0 ? s[i] : s[i + H2], sk = h0 ? s[i + H2] : s[i];
          const float qs = h0 ? q[i] : q[i + H2], qk = h0 ? q[i + H2] : q[i];
          s[i] = sk + __shfl_xor_sync(0xffffffffu, ss, 8);
          q[i] = qk + __shfl_xor_sync(0xffffffffu, qs, 8);
        }
        const int r0 = r1 + (h0 ? H2 : 0);
#pragma unroll
        for (int i = 0; i < H2; ++i) rmw(lane & 7, r0 + i, s[i], q[i]);
      } else {
        s[0] += __shfl_xor_sync(0xffffffffu, s[0], 8);
        q[0] += __shfl_xor_sync(0xffffffffu, q[0], 8);
        if (!h0) rmw(lane & 7, r1, s[0], q[0]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < H; ++i) rmw(lane & 15, r1 + i, s[i], q[i]);
    }
  } else {
    s[0] += __shfl_xor_sync(0xffffffffu, s[0], 16);
    q[0] += __shfl_xor_sync(0xffffffffu, q[0], 16);
    if constexpr (S8) {
      s[0] += __shfl_xor_sync(0xffffffffu, s[0], 8);
      q[0] += __shfl_xor_sync(0xffffffffu, q[0], 8);
    }
    if (lane < (S8 ? 8 : 16)) rmw(lane, 0, s[0], q[0]);
  }
}

// GroupNorm statistics of one 32-column unit held in v[] (bias not yet added): adds this thread's row to the per-warp
// partial sums of its sample (Sigma x, Sigma x^2 per group).  bias32 / qg: the unit's 32 bias values and the group of each
// 4-channel quad.  Rows beyond the layer's length contribute nothing (ok == false).
__device__ __forceinline__ void gn_accumulate_unit(const ConvJob& job, const float* v, const float* bias32, bool ok, int lane,
                                             float* pw, const unsigned char* qg) {
  constexpr int S = 8;
  float s[8], q[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const float4 bi = *reinterpret_cast<const float4*>(bias32 + r * 4);
    const float x0 = v[r * 4] + bi.x, x1 = v[r * 4 + 1] + bi.y, x2 = v[r * 4 + 2] + bi.z, x3 = v[r * 4 + 3] + bi.w;
    s[r] = ok ? (x0 + x1) + (x2 + x3) : 0.f;
    q[r] = ok ? fmaf(x0, x0, fmaf(x1, x1, fmaf(x2, x2, x3 * x3))) : 0.f;
  }
  if (job.cg_pow2) {
    // fold 4-channel quads of the same GroupNorm group inside the thread, then recursive halving across lanes
    if (job.cg < 8) {
      if (S == 8) gn_reduce_rows<8, true>(s, q, lane, pw, qg); else gn_reduce_rows<8, false>(s, q, lane, pw, qg);
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) { s[i] = s[2 * i] + s[2 * i + 1]; q[i] = q[2 * i] + q[2 * i + 1]; }
      if (job.cg < 16) {
        if (S == 8) gn_reduce_rows<4, true>(s, q, lane, pw, qg); else gn_reduce_rows<4, false>(s, q, lane, pw, qg);
      } else {
        s[0] += s[1]; q[0] += q[1]; s[1] = s[2] + s[3]; q[1] = q[2] + q[3];
        if (job.cg < 32) {
          if (S == 8) gn_reduce_rows<2, true>(s, q, lane, pw, qg); else gn_reduce_rows<2, false>(s, q, lane, pw, qg);
        } else {
          s[0] += s[1]; q[0] += q[1];
          if (S == 8) gn_reduce_rows<1, true>(s, q, lane, pw, qg); else gn_reduce_rows<1, false>(s, q, lane, pw, qg);
        }
      }
    }
  } else {
    // arbitrary group width (a multiple of 4): several quads may share a group, one lane adds them in order
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      s[r] += __shfl_xor_sync(0xffffffffu, s[r], 16);
      q[r] += __shfl_xor_sync(0xffffffffu, q[r], 16);
      if (S == 8) {
        s[r] += __shfl_xor_sync(0xffffffffu, s[r], 8);
        q[r] += __shfl_xor_sync(0xffffffffu, q[r], 8);
      }
      if (lane < S) {
        float2* pp = reinterpret_cast<float2*>(&pw[(lane * 8 + qg[r]) * 2]);
        float2 acc = *pp;
        acc.x += s[r];
        acc.y += q[r];
        *pp = acc;
      }
    }
  }
}

// GroupNorm statistics -> per-(sample, channel) scale / shift table.  Called by all epilogue threads once every warp has
// accumulated its partial sums: fixed-order combination over the warps (deterministic), then
//   y = x * sc + sf,  sc = rstd * gamma,  sf = (bias - mean) * sc + beta      (table rows of 2 * Nc + 4 floats in sh.part)
// L-split: the twin CTA (same samples, other half of the positions) holds the other half of every sum: the 128 combining
// threads write theirs into the twin's sh.xstat (st.shared::cluster + remote mbarrier arrive), wait for the twin's, and add
// the two in a fixed order (half 0 first), so both CTAs normalise with bit-identical statistics.
template <bool LS>
__device__ __forceinline__ void gn_finalize(const ConvJob& job, ConvShared& sh, int tid, int Nc, int lhalf = 0,
                                            int pair_base = 0, int rank = 0, uint32_t x_parity = 0) {
  constexpr int S = 8;
  named_bar_sync(1, kEpiThreads);
  if (tid < 128) {                       // one thread per (sample, group): fixed-order sum over the warps
    float a = 0.f, q2 = 0.f;
#pragma unroll
    for (int w = 0; w < kEpiWarps; ++w) {
      const float2 pv = *reinterpret_cast<const float2*>(&sh.part[(w * 128 + tid) * 2]);
      a += pv.x;
      q2 += pv.y;
    }
    if (LS && job.ls) {
      const uint32_t twin = (uint32_t)((pair_base ^ 2) + rank);
      st_remote_f32x2(mapa_u32(smem_u32(&sh.xstat[x_parity][tid][0]), twin), a, q2);
      mbar_arrive_remote(mapa_u32(smem_u32(&sh.xstat_full), twin));          // release.cluster: orders the store before it
      mbar_wait_cluster(&sh.xstat_full, x_parity, 0x530);
      const float2 o = *reinterpret_cast<const float2*>(&sh.xstat[x_parity][tid][0]);
      a = lhalf == 0 ? a + o.x : o.x + a;
      q2 = lhalf == 0 ? q2 + o.y : o.y + q2;
    }
    const float inv_n = 1.f / (float)(job.cg * job.gn_L);
    const float mean = a * inv_n;
    const float var = fmaxf(q2 * inv_n - mean * mean, 0.f);
    (&sh.stat[0][0][0])[tid * 2] = mean;
    (&sh.stat[0][0][0])[tid * 2 + 1] = rsqrtf(var + 1e-5f);
  }
  named_bar_sync(1, kEpiThreads);
  // scale/shift table: y = x * sc + sf with sc = rstd * gamma, sf = (bias - mean) * sc + beta
  const int tab_stride = 2 * Nc + 4;
  if (kEpiThreads % Nc == 0) {           // Nc in {32, 64, 128, ...}: no integer division in the loop
    const int c = tid % Nc;
    const int gi = sh.qgrp[c >> 2];
    const float ga = sh.par[1][c], be = sh.par[2][c], bi = sh.par[0][c];
    for (int bb = tid / Nc; bb < S; bb += kEpiThreads / Nc) {
      const float2 st = *reinterpret_cast<const float2*>(&sh.stat[bb][gi][0]);
      const float sc = st.y * ga;
      float2 o;
      o.x = sc;
      o.y = fmaf(bi - st.x, sc, be);
      *reinterpret_cast<float2*>(&sh.part[bb * tab_stride + 2 * c]) = o;
    }
  } else {
    for (int e = tid; e < S * Nc; e += kEpiThreads) {
      const int bb = e / Nc, c = e - bb * Nc;
      const int gi = sh.qgrp[c >> 2];
      const float2 st = *reinterpret_cast<const float2*>(&sh.stat[bb][gi][0]);
      const float sc = st.y * sh.par[1][c];
      float2 o;
      o.x = sc;
      o.y = fmaf(sh.par[0][c] - st.x, sc, sh.par[2][c]);
      *reinterpret_cast<float2*>(&sh.part[bb * tab_stride + 2 * c]) = o;
    }
  }
  named_bar_sync(1, kEpiThreads);
}

// Writes (optionally after adding the staged residual) 32 consecutive channels [cb, cb+32) of staged row `srow` as
// fp16 hi/lo planes: this thread's 64 B + 64 B, 16-byte slots XOR-swizzled like a TMA box (conflict-free).
__device__ __forceinline__ void staged_write(const ConvJob& job, uint32_t stg_data, int srow, int cb, float* v,
                                             bool add_res) {
  const int cw_shift = job.st_cw == 64 ? 6 : 5;
  const uint32_t row_bytes = (uint32_t)job.st_cw * 2;
  const uint32_t blk_stride = (uint32_t)job.st_blk_stride;
  const uint32_t plane_bytes = (uint32_t)job.st_nq * blk_stride;
  const uint32_t xr = job.st_cw == 64 ? (uint32_t)(srow & 7) : (uint32_t)((srow >> 1) & 3);
  const uint32_t i0 = (uint32_t)(cb & (job.st_cw - 1)) >> 3;
  const uint32_t base = stg_data + (uint32_t)(cb >> cw_shift) * blk_stride + (uint32_t)srow * row_bytes;
#pragma unroll
  for (int q4 = 0; q4 < 4; ++q4) {
    const uint32_t a_hi = base + (((i0 + q4) ^ xr) << 4);
    const uint32_t a_lo = a_hi + plane_bytes;
    if (add_res) {
      const uint4 hv = lds128(a_hi), lv = lds128(a_lo);
      const unsigned int* hh = reinterpret_cast<const unsigned int*>(&hv);
      const unsigned int* ll = reinterpret_cast<const unsigned int*>(&lv);
#pragma unroll
      for (int w = 0; w < 4; ++w) {   // (v + hi) + lo, one mixed-precision add each
        v[q4 * 8 + w * 2] = add_f32_h(add_f32_h(v[q4 * 8 + w * 2], (unsigned short)(hh[w] & 0xFFFFu)), (unsigned short)(ll[w] & 0xFFFFu));
        v[q4 * 8 + w * 2 + 1] = add_f32_h(add_f32_h(v[q4 * 8 + w * 2 + 1], (unsigned short)(hh[w] >> 16)), (unsigned short)(ll[w] >> 16));
      }
    }
    uint4 hv, lv;
    __half2* hh = reinterpret_cast<__half2*>(&hv);
    __half2* ll = reinterpret_cast<__half2*>(&lv);
#pragma unroll
    for (int w = 0; w < 4; ++w) split_half2(v[q4 * 8 + w * 2], v[q4 * 8 + w * 2 + 1], hh[w], ll[w]);
    sts128(a_hi, hv);
    sts128(a_lo, lv);
  }
}

// ================================ epilogue, TMEM rounds (horizons above 256) ================================
// The layer's l-tiles pass through TMEM job.lt_round at a time (mma_unit): after every round the accumulators are drained
// and handed back to the leader's MMA warps (sh.tile_ready, one arrival per CTA).  GroupNorm layers accumulate the
// statistics round by round and park the raw conv output in job.raw (fp32, written and later re-read by the SAME thread);
// after the last round the table is built and pass 2 streams the raw values back.  Bias-only layers (down / up-sampling,
// 1x1) finish each round directly.  Outputs go straight to global memory (these tiles exceed the staged epilogue).
// Only the RD = true kernel variants contain it: the H <= 256 kernels carry neither its registers nor its instructions.
template <bool TR>
__device__ __forceinline__ void epilogue_rounds(const ConvJob& job, int warp, int g, int b0, int n0, int rank,
                                             const float* tbias, int tbias_stride, ConvShared& sh, uint32_t tmem_base,
                                             uint32_t acc_count0, int pair_base = 0) {
  constexpr int S = 8;
  const int lane = threadIdx.x & 31, tid = threadIdx.x;
  const int Nc = job.Nc, mrows = job.mrows, ncols = job.ncols;
  const int cbase = mrows == 64 ? ((warp & 3) >> 1) * ncols : 0;
  const int row = (mrows == 64 ? (warp & 1) : (warp & 3)) * 32 + lane;
  const int slice = warp >> 2, b = row & (S - 1);
  const bool gn = job.epi == EPI_GN_MISH;
  const bool tb_uniform = tbias != nullptr && tbias_stride == 0;
  for (int i = tid; i < Nc; i += kEpiThreads) {
    sh.par[0][i] = __ldg(job.bias + n0 + i);
    sh.par[1][i] = gn ? __ldg(job.gamma + n0 + i) : 1.f;
    sh.par[2][i] = gn ? __ldg(job.beta + n0 + i) : 0.f;
    sh.par[3][i] = tb_uniform ? __ldg(tbias + n0 + i) : 0.f;
  }
  if (tid < (Nc >> 2)) sh.qgrp[tid] = (unsigned char)((tid * 4) / job.cg);
  if (gn)
    for (int i = tid; i < kEpiWarps * 256; i += kEpiThreads) sh.part[i] = 0.f;
  if (tid == 0) tma_store_wait_read();
  const uint32_t trow = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  const int nchunks = ncols >> 5;
  const int MTr = job.lt_round * job.Mt;              // tiles per phase held in TMEM during one round
  const int MT_all = job.n_ltiles * job.Mt;
  const int bg = g * kGroupSamples + b0 + b;
  float v[32];

  // one 32-channel piece of one output row: bias / normalise + Mish + time bias, + residual, stores
  auto finish = [&](int lrow, int cb) {
    if (gn) {
      const float* tab = &sh.part[b * (2 * Nc + 4) + 2 * cb];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 s01 = *reinterpret_cast<const float4*>(tab + r * 8);
        const float4 s23 = *reinterpret_cast<const float4*>(tab + r * 8 + 4);
        const float4 tb = *reinterpret_cast<const float4*>(&sh.par[3][cb + r * 4]);
        v[r * 4] = mish_f(fmaf(v[r * 4], s01.x, s01.y)) + tb.x;
        v[r * 4 + 1] = mish_f(fmaf(v[r * 4 + 1], s01.z, s01.w)) + tb.y;
        v[r * 4 + 2] = mish_f(fmaf(v[r * 4 + 2], s23.x, s23.y)) + tb.z;
        v[r * 4 + 3] = mish_f(fmaf(v[r * 4 + 3], s23.z, s23.w)) + tb.w;
      }
      if (tbias != nullptr && tbias_stride != 0) {
        const float4* tb = reinterpret_cast<const float4*>(tbias + (size_t)bg * tbias_stride + n0 + cb);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float4 t4 = __ldg(tb + e);
          v[e * 4] += t4.x; v[e * 4 + 1] += t4.y; v[e * 4 + 2] += t4.z; v[e * 4 + 3] += t4.w;
        }
      }
    } else {
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 bi = *reinterpret_cast<const float4*>(&sh.par[0][cb + r * 4]);
        v[r * 4] += bi.x; v[r * 4 + 1] += bi.y; v[r * 4 + 2] += bi.z; v[r * 4 + 3] += bi.w;
      }
    }
    const size_t rowoff = ((size_t)g * job.Lout + lrow) * kGroupSamples + b0 + b;
    if (job.res) {
      const uint4* rh = reinterpret_cast<const uint4*>(job.res + rowoff * job.res_C + n0 + cb);
      const uint4* rl = reinterpret_cast<const uint4*>(job.res + job.res_plane_stride + rowoff * job.res_C + n0 + cb);
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {
        const uint4 hv = __ldcg(rh + q4), lv = __ldcg(rl + q4);
        const __half2* hh = reinterpret_cast<const __half2*>(&hv);
        const __half2* ll = reinterpret_cast<const __half2*>(&lv);
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const float2 fh = __half22float2(hh[w]), fl = __half22float2(ll[w]);
          v[q4 * 8 + w * 2] += fh.x + fl.x;
          v[q4 * 8 + w * 2 + 1] += fh.y + fl.y;
        }
      }
    }
    if (job.out) {
      uint4* oh = reinterpret_cast<uint4*>(job.out + rowoff * job.out_C + n0 + cb);
      uint4* ol = reinterpret_cast<uint4*>(job.out + job.out_plane_stride + rowoff * job.out_C + n0 + cb);
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {
        uint4 hv, lv;
        __half2* hh = reinterpret_cast<__half2*>(&hv);
        __half2* ll = reinterpret_cast<__half2*>(&lv);
#pragma unroll
        for (int w = 0; w < 4; ++w) split_half2(v[q4 * 8 + w * 2], v[q4 * 8 + w * 2 + 1], hh[w], ll[w]);
        oh[q4] = hv;
        ol[q4] = lv;
      }
    }
    if (job.out_f32 && bg < job.B) {
      float* of = job.out_f32 + ((size_t)bg * job.Lout + lrow) * job.out_f32_C;
#pragma unroll
      for (int e = 0; e < 32; ++e)
        if (n0 + cb + e < job.Cout) of[n0 + cb + e] = v[e];
    }
  };

  for (int rd = 0; rd < job.n_rounds; ++rd) {
    if (warp == kPollWarp) mbar_wait(&sh.acc_full, (acc_count0 + rd) & 1, 0x520);
    named_bar_sync(1, kEpiThreads);      // (also orders the parameter loads above before their first use)
    tc_fence_after();
    const int t0 = rd * MTr;
    const int nt = min(MTr, MT_all - t0);
    if (gn) {
      // pass 1 of this round: statistics, and the raw accumulators (bias not yet added) out to the scratch
      for (int Tl = 0; Tl < nt; ++Tl) {
        const int j = ((t0 + Tl) * mrows + row) >> 3;
        const bool ok = j < job.Lj;
        for (int cc = (slice - Tl * nchunks) & 3; cc < nchunks; cc += 4) {
          tmem_ld32(trow + (uint32_t)(Tl * ncols + cc * 32), v);
          tmem_ld_wait();
          if (job.acc2) tmem_ld32_add(trow + (uint32_t)(job.acc2_off + Tl * ncols + cc * 32), v);
          const int cb = cbase + cc * 32;
          gn_accumulate_unit(job, v, &sh.par[0][cb], ok, lane, &sh.part[warp * 256], &sh.qgrp[cb >> 2]);
          if (ok) {
            float4* dst = reinterpret_cast<float4*>(job.raw + (((size_t)g * job.Lj + j) * kGroupSamples + b0 + b) * job.raw_C + n0 + cb);
#pragma unroll
            for (int e = 0; e < 8; ++e) dst[e] = make_float4(v[e * 4], v[e * 4 + 1], v[e * 4 + 2], v[e * 4 + 3]);
          }
        }
      }
    } else {
      // bias-only layers finish the round's tiles (both phases of a transposed conv) here
      for (int ph = 0; ph < job.nphase; ++ph)
        for (int Tl = 0; Tl < nt; ++Tl) {
          const int j = ((t0 + Tl) * mrows + row) >> 3;
          const int lrow = job.nphase == 2 ? 2 * j + ph : j;
          const int pt = ph * MTr + Tl;
          for (int cc = (slice - pt * nchunks) & 3; cc < nchunks; cc += 4) {
            tmem_ld32(trow + (uint32_t)(pt * ncols + cc * 32), v);
            tmem_ld_wait();
            if (job.acc2) tmem_ld32_add(trow + (uint32_t)(job.acc2_off + pt * ncols + cc * 32), v);
            if (j < job.Lj) finish(lrow, cbase + cc * 32);
          }
        }
    }
    if (rd + 1 < job.n_rounds) {           // give the TMEM columns back: the next round's MMAs overwrite them in both CTAs
      tc_fence_before();
      named_bar_sync(1, kEpiThreads);
      if (tid == 0) {
        if (rank == 0) mbar_arrive(&sh.tile_ready);
        else mbar_arrive_remote(mapa_u32(smem_u32(&sh.tile_ready), (uint32_t)pair_base));
      }
    }
  }
  if (gn) {
    gn_finalize<false>(job, sh, tid, Nc);   // (starts with a barrier: every warp's partial sums are complete)
    // pass 2: every thread streams back the raw values it parked (same thread, same addresses: no fence needed)
    for (int T = 0; T < MT_all; ++T) {
      const int j = (T * mrows + row) >> 3;
      if (j >= job.Lj) continue;
      const int Tl = T % MTr;               // the unit -> warp assignment used when the tile was drained
      for (int cc = (slice - Tl * nchunks) & 3; cc < nchunks; cc += 4) {
        const int cb = cbase + cc * 32;
        const float4* src = reinterpret_cast<const float4*>(job.raw + (((size_t)g * job.Lj + j) * kGroupSamples + b0 + b) * job.raw_C + n0 + cb);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float4 t4 = __ldcg(src + e);
          v[e * 4] = t4.x; v[e * 4 + 1] = t4.y; v[e * 4 + 2] = t4.z; v[e * 4 + 3] = t4.w;
        }
        finish(j, cb);
      }
    }
  }
}

// ================================ epilogue (warps 0..15) ===================================
// Warp w reads TMEM lanes 32*(w%4).. (hardware rule) and takes the (tile, 32-column chunk) units u with
// u % 4 == w / 4.  Thread <-> GEMM row (l, b); b = row % S is constant per thread across tiles.
// RD: the "big tile" kernel variants (see mma_unit) also carry the wave-staged pass 2.
template <bool TR, bool RD, bool LS, bool MC>
__device__ __forceinline__ void epilogue_unit(const ConvJob& job, const ConvMaps* maps, int warp, int g, int b0, int n0,
                                              int rank, const float* tbias, int tbias_stride, uint32_t stg, ConvShared& sh,
                                              uint32_t tmem_base, RoleState& rs, unsigned long long* tr) {
  const int lane = threadIdx.x & 31;
  const int tid = threadIdx.x;                 // 0..511
  if (tid != 0) tr = nullptr;
  const int S = job.S, Nc = job.Nc;
  const int s_shift = 3;                         // S == 8
  const int mrows = job.mrows, ncols = job.ncols;
  // mrows 128: TMEM lane == row of the CTA's 128-row tile, all Nc channels in columns [0, Nc).  mrows 64 (M = 128 pair
  // MMA): lanes 0-63 hold the CTA's 64 rows for channels [0, Nc/2), lanes 64-127 the same rows for channels [Nc/2, Nc).
  const int cbase = mrows == 64 ? ((warp & 3) >> 1) * ncols : 0;   // first channel (within the unit) this warp's lanes hold
  const int MT = job.n_ltiles * job.Mt;
  const int row = (mrows == 64 ? (warp & 1) : (warp & 3)) * 32 + lane;      // row inside the CTA's tile
  const int slice = warp >> 2;
  const int b = row & (S - 1);
  const bool gn = job.epi == EPI_GN_MISH;
  const bool tb_uniform = tbias != nullptr && tbias_stride == 0;
  const int loff = LS ? rs.lhalf * job.ls_out : 0;   // L-split: first output position of this pair's half (0 otherwise)

  for (int i = tid; i < Nc; i += kEpiThreads) {
    sh.par[0][i] = __ldg(job.bias + n0 + i);
    sh.par[1][i] = gn ? __ldg(job.gamma + n0 + i) : 1.f;
    sh.par[2][i] = gn ? __ldg(job.beta + n0 + i) : 0.f;
    sh.par[3][i] = tb_uniform ? __ldg(tbias + n0 + i) : 0.f;
    if (job.aux) sh.par[4][i] = __ldg(job.aux_bias + n0 + i);
  }
  if (tid < (Nc >> 2)) sh.qgrp[tid] = (unsigned char)((tid * 4) / job.cg);
  if (gn) {
    for (int i = tid; i < kEpiWarps * 256; i += kEpiThreads) sh.part[i] = 0.f;
  }
  trace_mark<TR>(tr, 8);
  // The previous layer's TMA store may still be reading the staging buffer (deferred wait of a resident link).
  if (tid == 0) tma_store_wait_read();
  // one warp waits for the accumulator; everybody else parks on the hardware barrier (no issue slots burnt)
  if (warp == kPollWarp) mbar_wait(&sh.acc_full, rs.acc_count & 1, 0x500);
  rs.acc_count++;
  named_bar_sync(1, kEpiThreads);
  tc_fence_after();
  trace_mark<TR>(tr, 9);
  // Staged epilogue: every MMA has completed, so the operand stages are dead and the unit's output tile is
  // assembled there.  The residual tile (temporal.py:46) is TMA-loaded into the same buffer now; it lands
  // while pass 1 runs.  Block (plane p, channel block q) = kStageGap zero bytes | st_rows rows | kStageGap zero
  // bytes at stg + (p * st_nq + q) * st_blk_stride; the gaps are the conv halo of a resident consumer.
  const bool wave = RD && job.st_wave != 0;    // big tiles: staged and stored wave by wave (see pass 2)
  const bool staged = job.st_cw != 0 && !wave;
  const bool res_staged = staged && job.res != nullptr;
  const uint32_t blk_bytes = (uint32_t)(job.st_rows * job.st_cw * 2);
  const uint32_t blk_stride = (uint32_t)job.st_blk_stride;
  const uint32_t stg_data = stg + kStageGap;
  if (res_staged) {
    if (tid == 0) {
      tma_store_wait_all();   // the residual may be an auxiliary tile this thread stored one layer ago without waiting
      mbar_arrive_expect_tx(&sh.res_full, 2u * job.st_nq * blk_bytes);
      for (int p = 0; p < 2; ++p)
        for (int q = 0; q < job.st_nq; ++q)
          tma_load_5d(stg_data + (uint32_t)(p * job.st_nq + q) * blk_stride, &maps->r_map, &sh.res_full,
                      n0 + q * job.st_cw, b0, loff, g, p);
    }
    __syncwarp();
  }
  const uint32_t trow = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  const int nchunks = ncols >> 5;               // 32-column units per tile in this warp's lanes
  float v[32];

  if (job.aux) {
    // ---- auxiliary 1x1 conv (the block's residual_conv): bias only, staged and stored first; its TMA store
    //      drains while pass 1 runs, and the staging buffer is reclaimed before pass 2 writes it ----
    for (int T = 0; T < MT; ++T) {
      const int jrow = (T * mrows + row) >> s_shift;
      for (int cc = (slice - T * nchunks) & 3; cc < nchunks; cc += 4) {
        tmem_ld32(trow + (uint32_t)(job.aux_col0 + T * ncols + cc * 32), v);
        tmem_ld_wait();
        if (jrow >= job.Lj) continue;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const float4 bi = *reinterpret_cast<const float4*>(&sh.par[4][cbase + cc * 32 + r * 4]);
          v[r * 4] += bi.x; v[r * 4 + 1] += bi.y; v[r * 4 + 2] += bi.z; v[r * 4 + 3] += bi.w;
        }
        staged_write(job, stg_data, jrow * S + b, cbase + cc * 32, v, false);
      }
    }
    fence_proxy_async_smem();
    named_bar_sync(1, kEpiThreads);
    if (tid == 0) {
      for (int p = 0; p < 2; ++p)
        for (int q = 0; q < job.st_nq; ++q)
          tma_store_5d(&maps->p_map, stg_data + (uint32_t)(p * job.st_nq + q) * blk_stride, n0 + q * job.st_cw, b0, loff,
                       g, p);
      tma_store_commit();
    }
    __syncwarp();
  }

  if (gn) {
    // ---- pass 1: GroupNorm statistics over (channels of group) x (all L positions), per sample ----
    for (int T = 0; T < MT; ++T) {
      const bool ok = ((T * mrows + row) >> s_shift) < job.Lj;
      for (int cc = (slice - T * nchunks) & 3; cc < nchunks; cc += 4) {   // units u = T * nchunks + cc with u % 4 == slice
        tmem_ld32(trow + (uint32_t)(T * ncols + cc * 32), v);
        tmem_ld_wait();
        if (job.acc2) tmem_ld32_add(trow + (uint32_t)(job.acc2_off + T * ncols + cc * 32), v);
        // the two issuer warps' accumulator sets are folded ONCE: the sum goes back into set 0 (tcgen05.st), and pass 2
        // (same thread, same unit) needs a single TMEM load
        if (job.acc2) tmem_st32(trow + (uint32_t)(T * ncols + cc * 32), v);
        gn_accumulate_unit(job, v, &sh.par[0][cbase + cc * 32], ok, lane, &sh.part[warp * 256], &sh.qgrp[(cbase >> 2) + cc * 8]);
      }
    }
    if (job.acc2) tmem_st_wait();
    trace_mark<TR>(tr, 40);
    if (job.aux && tid == 0) tma_store_wait_read();   // the auxiliary tile has left the staging buffer
    gn_finalize<LS>(job, sh, tid, Nc, rs.lhalf, rs.pair_base, rank, rs.x_count & 1);
    if (LS && job.ls) rs.x_count++;
  }
  trace_mark<TR>(tr, 10);

  // ---- pass 2: normalise, Mish, + time bias, + residual, split to fp16 hi/lo, store ----
  const int bg = g * kGroupSamples + b0 + b;
  const int PT = job.nphase * MT;
  if (res_staged) {
    mbar_wait(&sh.res_full, rs.res_count & 1, 0x600);
    rs.res_count++;
  }
  // normalise + Mish + time bias (GroupNorm layers) or bias (others) of the 32 channels [cb, cb + 32) held in v[]
  auto norm_unit = [&](int cb) {
    if (gn) {
      const float* trow_tab = &sh.part[b * (2 * Nc + 4) + 2 * cb];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 s01 = *reinterpret_cast<const float4*>(trow_tab + r * 8);
        const float4 s23 = *reinterpret_cast<const float4*>(trow_tab + r * 8 + 4);
        const float4 tb = *reinterpret_cast<const float4*>(&sh.par[3][cb + r * 4]);
        v[r * 4] = mish_f(fmaf(v[r * 4], s01.x, s01.y)) + tb.x;
        v[r * 4 + 1] = mish_f(fmaf(v[r * 4 + 1], s01.z, s01.w)) + tb.y;
        v[r * 4 + 2] = mish_f(fmaf(v[r * 4 + 2], s23.x, s23.y)) + tb.z;
        v[r * 4 + 3] = mish_f(fmaf(v[r * 4 + 3], s23.z, s23.w)) + tb.w;
      }
    } else {
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 bi = *reinterpret_cast<const float4*>(&sh.par[0][cb + r * 4]);
        v[r * 4] += bi.x; v[r * 4 + 1] += bi.y; v[r * 4 + 2] += bi.z; v[r * 4 + 3] += bi.w;
      }
    }
    if (gn && tbias != nullptr && tbias_stride != 0) {
      const float4* tb = reinterpret_cast<const float4*>(tbias + (size_t)bg * tbias_stride + n0 + cb);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float4 t4 = __ldg(tb + e);
        v[e * 4] += t4.x; v[e * 4 + 1] += t4.y; v[e * 4 + 2] += t4.z; v[e * 4 + 3] += t4.w;
      }
    }
  };
  if (RD && wave) {
    // ---- big tiles (they exceed the staging budget as a whole): the (tile, 32-channel) units go out in WAVES of four,
    //      one unit per slice group = 512 rows x 32 | 256 x 64 | 128 x 128 channels = 64 KB of hi/lo planes, assembled in one
    //      of two staging buffers in the TMA box layout and written with TMA stores while the next wave is computed.
    //      The residual arrives the same way: the wave's box is TMA-loaded into the buffer before the wave is computed. ----
    const uint32_t wave_bytes = 2u * (uint32_t)job.st_nq * blk_stride;
    const int n_units = MT * nchunks;
    const int n_waves = (n_units + 3) >> 2;
    const int tiles_per_wave = 4 / nchunks;          // nchunks in {1, 2, 4}
    auto load_res = [&](int w) {                      // tid 0: residual boxes of wave w into buffer w & 1
      const uint32_t buf = stg + (uint32_t)(w & 1) * wave_bytes;
      mbar_arrive_expect_tx(&sh.res_full, wave_bytes);
      for (int p = 0; p < 2; ++p)
        for (int q = 0; q < job.st_nq; ++q)
          tma_load_5d(buf + (uint32_t)(p * job.st_nq + q) * blk_stride, &maps->r_map, &sh.res_full, n0 + q * job.st_cw, b0,
                      loff + w * tiles_per_wave * (mrows >> 3), g, p);
    };
    const bool wres = job.res != nullptr;
    if (wres && tid == 0) { tma_store_wait_all(); load_res(0); }
    for (int w = 0; w < n_waves; ++w) {
      const uint32_t buf = stg + (uint32_t)(w & 1) * wave_bytes;
      const int u = 4 * w + slice;
      const int T = u / nchunks, cc = u - T * nchunks;
      const int T0 = w * tiles_per_wave;
      const int j = (T * mrows + row) >> s_shift;
      if (wres) {
        mbar_wait(&sh.res_full, rs.res_count & 1, 0x610);
        rs.res_count++;
      }
      if (u < n_units) {
        const int cb = cbase + cc * 32;
        tmem_ld32(trow + (uint32_t)(T * ncols + cc * 32), v);
        tmem_ld_wait();
        if (job.acc2 && !gn) tmem_ld32_add(trow + (uint32_t)(job.acc2_off + T * ncols + cc * 32), v);
        if (j < job.Lj) {
          norm_unit(cb);
          staged_write(job, buf, (T - T0) * mrows + row, cb, v, wres);
        }
      }
      fence_proxy_async_smem();
      named_bar_sync(1, kEpiThreads);
      if (tid == 0) {
        for (int p = 0; p < 2; ++p)
          for (int q = 0; q < job.st_nq; ++q)
            tma_store_5d(&maps->o_map, buf + (uint32_t)(p * job.st_nq + q) * blk_stride, n0 + q * job.st_cw, b0,
                         loff + T0 * (mrows >> 3), g, p);
        tma_store_commit();
        if (w + 1 < n_waves) {
          // buffer (w + 1) & 1 was last read by the store of wave w - 1: at most this wave's group may still be reading
          asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          if (wres) load_res(w + 1);
        }
      }
      if (w + 1 < n_waves && !wres) named_bar_sync(1, kEpiThreads);   // (with a residual the res_full wait orders the reuse)
    }
    if (tid == 0) tma_store_wait_all();   // consumers TMA-load this tile from global memory
    __syncwarp();
  } else
  for (int pt = 0; pt < PT; ++pt) {
    const int ph = pt >= MT ? 1 : 0;
    const int T = pt - ph * MT;
    const int j = (T * mrows + row) >> s_shift;
    const bool ok = j < job.Lj;
    const int lrow = job.nphase == 2 ? 2 * j + ph : j;
    const size_t rowoff = ((size_t)g * job.Lfull_out + loff + lrow) * kGroupSamples + b0 + b;
    for (int cc = (slice - pt * nchunks) & 3; cc < nchunks; cc += 4) {
      const int cb = cbase + cc * 32;
      tmem_ld32(trow + (uint32_t)(pt * ncols + cc * 32), v);
      tmem_ld_wait();
      if (job.acc2 && !gn) tmem_ld32_add(trow + (uint32_t)(job.acc2_off + pt * ncols + cc * 32), v);   // (GroupNorm layers: folded in pass 1)
      if (!ok) continue;
      norm_unit(cb);
      if (staged) {
        staged_write(job, stg_data, lrow * S + b, cb, v, res_staged);
      } else {
      if (job.res) {
        // (L2-coherent loads: inside the persistent kernel this data was written by another SM moments ago)
        const uint4* rh = reinterpret_cast<const uint4*>(job.res + rowoff * job.res_C + n0 + cb);
        const uint4* rl = reinterpret_cast<const uint4*>(job.res + job.res_plane_stride + rowoff * job.res_C + n0 + cb);
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const uint4 hv = __ldcg(rh + q4), lv = __ldcg(rl + q4);
          const __half2* hh = reinterpret_cast<const __half2*>(&hv);
          const __half2* ll = reinterpret_cast<const __half2*>(&lv);
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            const float2 fh = __half22float2(hh[w]), fl = __half22float2(ll[w]);
            v[q4 * 8 + w * 2] += fh.x + fl.x;
            v[q4 * 8 + w * 2 + 1] += fh.y + fl.y;
          }
        }
      }
      if (job.out) {
        uint4* oh = reinterpret_cast<uint4*>(job.out + rowoff * job.out_C + n0 + cb);
        uint4* ol = reinterpret_cast<uint4*>(job.out + job.out_plane_stride + rowoff * job.out_C + n0 + cb);
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          uint4 hv, lv;
          __half2* hh = reinterpret_cast<__half2*>(&hv);
          __half2* ll = reinterpret_cast<__half2*>(&lv);
#pragma unroll
          for (int w = 0; w < 4; ++w) split_half2(v[q4 * 8 + w * 2], v[q4 * 8 + w * 2 + 1], hh[w], ll[w]);
          oh[q4] = hv;
          ol[q4] = lv;
        }
      }
      }
      if (job.out_f32 && job.f32_stride) {
        // fp32 public-layout output, staged: one sample = Lout * C contiguous floats (+4 floats of bank skew per
        // sample: lanes are 4 positions x 8 samples, 35-float rows -> conflict-free), copied out coalesced below
        const uint32_t so = stg + 4u * (uint32_t)(b * job.f32_stride + lrow * job.out_f32_C + n0 + cb);
#pragma unroll
        for (int e = 0; e < 32; ++e)
          if (n0 + cb + e < job.Cout) sts32(so + 4u * e, v[e]);
      } else if (job.out_f32 && bg < job.B) {
        float* of = job.out_f32 + ((size_t)bg * job.Lfull_out + loff + lrow) * job.out_f32_C;
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const int c = n0 + cb + e;
          if (c < job.Cout) of[c] = v[e];
        }
      }
    }
  }
  if (job.out_f32 && job.f32_stride) {
    named_bar_sync(1, kEpiThreads);
    const int per = (job.Lout * job.out_f32_C) >> 2;          // float4 per sample
    for (int i = tid; i < S * per; i += kEpiThreads) {
      const int bb = i / per, o = i - bb * per;
      const int sg = g * kGroupSamples + b0 + bb;
      if (sg < job.B)
        reinterpret_cast<float4*>(job.out_f32 + ((size_t)sg * job.Lfull_out + loff) * job.out_f32_C)[o] =
            lds128f(stg + 4u * (uint32_t)(bb * job.f32_stride) + 16u * (uint32_t)o);
    }
  }
  trace_mark<TR>(tr, 14);
  if (staged) {
    if (job.next_res && !job.gaps_clean) {
      // zero the halo gaps of every block (other layers' operand stages alias this memory)
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);
      const int per_blk = 2 * kStageGap / 16;
      for (int i = tid; i < 2 * job.st_nq * per_blk; i += kEpiThreads) {
        const int blk = i / per_blk, o = i - blk * per_blk;
        const bool leading = o < kStageGap / 16;
        // L-split: the gap towards the other half holds ITS edge rows (written by the twin CTA, below): only the outer one is zero
        if (LS && job.ls && leading == (rs.lhalf != 0)) continue;
        const uint32_t off = leading ? (uint32_t)o * 16u : kStageGap + blk_bytes + (uint32_t)(o - kStageGap / 16) * 16u;
        sts128(stg + (uint32_t)blk * blk_stride + off, z);
      }
    }
    if (LS && job.ls && job.next_res) {
      // The consumer's conv halo across the split: this CTA's two edge positions (16 rows of every block, raw bytes: the rows
      // keep their swizzle phase) go into the twin CTA's gap, once the twin's MMAs of THIS layer no longer read the rows that
      // are there (halo_free).  The twin's arrival on our leader's tile_ready (below) covers its copy into our gaps.
      fence_proxy_async_smem();
      if (warp == kPollWarp) mbar_wait_cluster(&sh.halo_free, rs.h_count & 1, 0x540);
      named_bar_sync(1, kEpiThreads);            // every row of this tile is staged, and the twin's gap may be written
      const uint32_t rb = (uint32_t)job.st_cw * 2u;                                     // bytes per staged row
      const uint32_t src0 = rs.lhalf == 0 ? kStageGap + blk_bytes - 16u * rb : kStageGap;          // last / first two positions
      const uint32_t dst0 = rs.lhalf == 0 ? kStageGap - 16u * rb : kStageGap + blk_bytes;          // twin's leading / trailing gap
      const uint32_t twin = (uint32_t)((rs.pair_base ^ 2) + rank);
      const int per_blk16 = (int)(16u * rb / 16u);
      for (int i = tid; i < 2 * job.st_nq * per_blk16; i += kEpiThreads) {
        const int blk = i / per_blk16, o = i - blk * per_blk16;
        const uint4 v = lds128(stg + (uint32_t)blk * blk_stride + src0 + 16u * (uint32_t)o);
        st_remote_b128(mapa_u32(stg + (uint32_t)blk * blk_stride + dst0 + 16u * (uint32_t)o, twin), v);
      }
      fence_proxy_async_all();                   // remote generic-proxy writes -> the twin's tensor-core reads
    }
    fence_proxy_async_smem();             // generic-proxy smem writes -> visible to the TMA store engine / next MMAs
    tc_fence_before();                    // this thread's TMEM reads are complete before the hand-over below
    named_bar_sync(1, kEpiThreads);
    if (job.next_res && threadIdx.x == 0) {   // releases the leader's MMA warps waiting in the next layer (one arrival per CTA)
      if (rank == 0) mbar_arrive(&sh.tile_ready);
      else mbar_arrive_remote(mapa_u32(smem_u32(&sh.tile_ready), (uint32_t)((LS || MC) ? rs.pair_base : 0)));
      if (LS && job.ls) mbar_arrive_remote(mapa_u32(smem_u32(&sh.tile_ready), (uint32_t)(rs.pair_base ^ 2)));   // ... and the twin pair's
    }
    trace_mark<TR>(tr, 15);
    if (tid == 0 && !job.no_store) {
      for (int p = 0; p < 2; ++p)
        for (int q = 0; q < job.st_nq; ++q)
          tma_store_5d(&maps->o_map, stg_data + (uint32_t)(p * job.st_nq + q) * blk_stride, n0 + q * job.st_cw, b0, loff,
                       g, p);
      tma_store_commit();
      // a non-resident consumer TMA-loads this tile from global, and the staging is reused by the next unit's
      // operand loads: the store must be complete.  A resident consumer reads shared memory; the wait moves
      // to its own epilogue (tma_store_wait_read above).
      if (!job.next_res) tma_store_wait_all();
    }
    __syncwarp();
  }
  if (LS && job.ls) rs.h_count++;   // (one halo_free phase per unit, waited for or not)
  trace_mark<TR>(tr, 11);
}

// One work unit = (sample sub-group, N slice) of one layer, executed by all roles of the CTA.
// rank = this CTA's rank in its pair: it owns samples b0 = 8 * rank .. of group g.
template <bool TR, bool RD, bool LS, bool MC = false>
__device__ __forceinline__ void run_unit(const ConvJob& job, const ConvMaps* maps, int warp, int rank,
                                         int g, int n0, const float* tbias, int tbias_stride, uint32_t smem,
                                         ConvShared& sh, uint32_t tmem_base, RoleState& rs,
                                         unsigned long long* tr = nullptr, const ConvJob* nj = nullptr,
                                         const CUtensorMap* n_w_map = nullptr, int n_n0 = 0,
                                         const CUtensorMap* n_wq_map = nullptr) {
  const uint32_t a_smem = smem;
  const uint32_t w_smem = smem + (job.a_res ? job.w_off_res : job.n_a_stages * job.a_stage_bytes);
  const int b0 = rank * 8;
  if (warp == kProducerWarp) {
    producer_unit<TR, LS, MC>(job, maps, &maps->a_map[0], &maps->w_map, &maps->x_map, g, b0, n0, rank, a_smem, w_smem, sh, rs, tr, smem, nj,
                              n_w_map, n_n0, n_wq_map);
  } else if (warp == kMmaWarp || warp == kMmaWarpB) {
    if (rank == 0) mma_unit<TR, RD, LS, MC>(job, smem, a_smem, w_smem, sh, tmem_base, rs, tr, warp == kMmaWarp ? 0 : 1);   // the leader issues for the pair
  } else if (RD && job.n_rounds > 1) {
    epilogue_rounds<TR>(job, warp, g, b0, n0, rank, tbias, tbias_stride, sh, tmem_base, rs.acc_count, MC ? rs.pair_base : 0);
    rs.acc_count += job.n_rounds;
  } else {
    epilogue_unit<TR, RD, LS, MC>(job, maps, warp, g, b0, n0, rank, tbias, tbias_stride, smem, sh, tmem_base, rs, tr);
  }
}

// Pair set-up / tear-down shared by both kernels: barriers of both CTAs exist before any remote arrival or pair TMA can
// target them, and nobody frees TMEM or leaves while the leader's MMAs may still read the peer's shared memory.
__device__ __forceinline__ uint32_t pair_begin(ConvShared& sh, int warp, uint32_t tmem_cols, bool mc = false, bool ls = false) {
  if (threadIdx.x == 0) init_shared(sh, mc, ls);
  __syncthreads();
  if (warp == kMmaWarp) tmem_alloc_pair(&sh.tmem_slot, tmem_cols);
  tc_fence_before();
  __syncwarp();
  cluster_sync_all();
  tc_fence_after();
  return sh.tmem_slot;
}
__device__ __forceinline__ void pair_end(int warp, uint32_t tmem_base, uint32_t tmem_cols) {
  tc_fence_before();
  __syncwarp();
  cluster_sync_all();
  if (warp == kMmaWarp) tmem_dealloc_pair(tmem_base, tmem_cols);
}

// 1024-byte aligned start of the dynamic shared memory, as a shared-window address (warp-uniform)
__device__ __forceinline__ uint32_t aligned_dyn_smem() {
  extern __shared__ uint8_t smem_raw[];
  return (smem_u32(smem_raw) + 1023u) & ~1023u;
}

// ------------------------------------------------------------------------------------------------
// per-layer launch: grid = (2 CTAs per 16-sample group, N slices), clusters of 2 along x (the CTA pair)
// ------------------------------------------------------------------------------------------------
// LS: L-split launch (conv_job.h): clusters of 4 = two pairs per group, pair h = cluster rank / 2 owns half h of the positions.
template <bool RD, bool LS>
__global__ void __cluster_dims__(LS ? 4 : 2, 1, 1) __launch_bounds__(kConvThreads, 1)
conv_gemm_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ ConvJob job) {
  const uint32_t smem = aligned_dyn_smem();
  __shared__ ConvShared sh;
  const int warp = warp_idx_sync();
  const int crank = (int)cluster_ctarank();
  const int rank = crank & 1;
  if (warp == kProducerWarp && elect_one_sync()) {
    tma_prefetch_desc(&maps.a_map[0]);
    if (job.nsrc > 1) tma_prefetch_desc(&maps.a_map[1]);
    tma_prefetch_desc(&maps.w_map);
  }
  const uint32_t tmem_base = pair_begin(sh, warp, job.tmem_cols);
  RoleState rs;
  rs.lhalf = LS ? crank >> 1 : 0;
  rs.pair_base = crank & ~1;
  run_unit<false, RD, LS>(job, &maps, warp, rank, (int)(blockIdx.x >> (LS ? 2 : 1)), blockIdx.y * job.Nc, job.tbias, job.tbias_stride,
                      smem, sh, tmem_base, rs);
  pair_end(warp, tmem_base, job.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// persistent sampling loop: cluster c of 2 CTAs owns samples [16c, 16c+16)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void job_sync(bool wrote_global, bool issues_tma) {
  if (wrote_global) fence_proxy_async_global();   // generic-proxy stores -> visible to later TMA (async proxy) loads
  tc_fence_before();
  __syncwarp();
  cluster_sync_all();
  tc_fence_after();
  if (issues_tma) fence_proxy_async_global();
}

// Same hand-over when producer and consumer are the SAME CTA (S = 8 layers: each CTA owns its 8 samples end to
// end): a block barrier is enough, the peer CTA is not involved.
__device__ __forceinline__ void local_sync(bool wrote_global, bool issues_tma) {
  if (wrote_global) fence_proxy_async_global();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (issues_tma) fence_proxy_async_global();
}

// The attention stages of the AT kernel variants.  (With one lane per softmax row the core stage needed more registers than
// the 96 this kernel has and was kept out of line; with its 4 x 8 register block it inlines with 148 bytes of spills, and the
// pass is 4 % faster than through the call.)
__device__ __forceinline__ void attn_stage_call(const ConvJob& job, int g, int b0, float* sm) {
  if (job.kind == JOB_ATTN_NORM) attn_norm_stage(job, g, b0, threadIdx.x);
  else attn_core_stage(job, g, b0, threadIdx.x, sm);
}

// AT: the layer list carries attention stages (attn_stages.cuh); a variant of its own so that the common kernels keep their
// register allocation.
// MC: clusters of four = two pairs working on two GROUPS (2c, 2c + 1) that share every weight tile by TMA multicast (see
// producer_unit); layers hand over through the cluster barrier wherever they are not resident links.
template <bool TR, bool RD, bool LS, bool AT, bool MC = false>
__global__ void __cluster_dims__((LS || MC) ? 4 : 2, 1, 1) __launch_bounds__(kConvThreads, 1)
sample_loop_kernel(const __grid_constant__ LoopParams lp) {
  static_assert(!(MC && (AT || LS)), "weight multicast exists for the unsplit kernels without attention stages only");
  const uint32_t smem = aligned_dyn_smem();
  __shared__ ConvShared sh;
  __shared__ StepParams sstep;
  const int warp = warp_idx_sync();
  const int crank = (int)cluster_ctarank();
  const int rank = crank & 1;
  const int g = MC ? 2 * (int)cluster_id_x() + (crank >> 1) : (int)cluster_id_x();
  const uint32_t tmem_base = pair_begin(sh, warp, 512, MC, LS);
  RoleState rs;
  rs.lhalf = LS ? crank >> 1 : 0;
  rs.pair_base = crank & ~1;

  // K4 works on whole samples: 8 per CTA, or 4 per CTA of an L-split cluster (every CTA of the group takes a share)
  const long long ht4 = ((long long)lp.H * lp.T) >> 2;           // float4 per sample
  const int k4_n = LS ? 4 : 8;
  const int s_lo = g * kGroupSamples + (MC ? rank : crank) * k4_n;
  const int s_hi = min(lp.B, s_lo + k4_n);

  for (int k = 0; k <= lp.n_steps; ++k) {
    if (k > 0) {
      const int t_index = lp.steps[k].t_index;
      for (int ji = 0; ji < lp.n_jobs; ++ji) {
        const ConvJob& job = lp.jobs[ji];   // kernel-parameter (constant) space: warp-uniform
        unsigned long long* tr = (TR && lp.trace && g == lp.trace_cluster && (MC ? rank : crank) == 0 && k == lp.trace_step)
                                     ? lp.trace + (size_t)ji * kTraceSlots : nullptr;
        if (TR && threadIdx.x == 0 && tr) { trace_mark<TR>(tr, 0); tr[1] = (unsigned long long)clock64(); }   // slot 1: SM clock (effective MHz)
        if (TR && threadIdx.x == kMmaWarp * 32 && tr) trace_mark<TR>(tr, 22);
#ifdef DM_DEBUG_WAIT
        if (threadIdx.x == 0) g_dm_dbg_pos[blockIdx.x] = (unsigned)ji | ((unsigned)k << 7);
#endif
        if (AT && job.kind != JOB_CONV) {
          // CUDA-core stage of an attention block: the epilogue warps work on this CTA's 8 samples, global -> global
          if (warp < kEpiWarps) {
            extern __shared__ uint8_t smem_raw[];
            attn_stage_call(job, g, rank * 8, reinterpret_cast<float*>(smem_raw + (smem - smem_u32(smem_raw))));
            fence_proxy_async_smem();   // the next layer's TMA loads overwrite this shared memory
          }
          if (threadIdx.x == 0) trace_mark<TR>(tr, 12);
          local_sync(warp < kEpiWarps, warp == kProducerWarp);
          if (threadIdx.x == 0) trace_mark<TR>(tr, 13);
          continue;
        }
        const int slices = job.slices;
        const int off = job.tb_off;
        const float* tb = off >= 0 ? lp.tb_table + (size_t)t_index * lp.TB + off : nullptr;
        const ConvMaps* maps = &lp.maps[ji];
        if (warp == kProducerWarp && ji + 1 < lp.n_jobs && elect_one_sync()) {   // hide next layer's descriptor fetch
          tma_prefetch_desc(&lp.maps[ji + 1].a_map[0]);
          tma_prefetch_desc(&lp.maps[ji + 1].w_map);
        }
        // the unit after this job's last one: first unit of the next job (next step's first job after the last)
        const int nji = ji + 1 < lp.n_jobs ? ji + 1 : 0;
        const ConvJob& njob = lp.jobs[nji];
        const bool has_next = ji + 1 < lp.n_jobs || k < lp.n_steps;
        for (int sl = 0; sl < slices; ++sl) {      // N slices of the layer, one after the other, by the whole pair
          const bool last_unit = sl + 1 >= slices;
          if (TR && threadIdx.x == kMmaWarp * 32 && tr && sl == 0) trace_mark<TR>(tr, 23);
          run_unit<TR, RD, LS, MC>(job, maps, warp, rank, g, sl * job.Nc, tb, 0, smem, sh, tmem_base, rs, sl == 0 ? tr : nullptr,
                                   last_unit && has_next ? &njob : nullptr, &lp.maps[nji].w_map, 0, &lp.maps[nji].wq_map);
          if (sl + 1 < slices) {   // TMEM / staging smem are reused by the next unit (the peer's side of the pair is
            tc_fence_before();     // protected through the full barriers: its loads for the next unit follow its barrier)
            if (LS || MC) cluster_sync_all();   // (L-split: also keeps the twin's next statistics exchange one use behind at most)
            else __syncthreads();
            tc_fence_after();
          }
        }
        if (threadIdx.x == 0) trace_mark<TR>(tr, 12);
        if (job.next_res) {
          // resident link: NO barrier.  The consumer is this CTA's next layer; its MMA warps wait on sh.tile_ready, its
          // producer waited for this layer's accumulator, its epilogue warps are the threads that just finished.
#ifdef DM_NO_RUNAHEAD   // A/B aid: the old block-wide hand-over
          tc_fence_before();
          __syncthreads();
          tc_fence_after();
#endif
        } else if (job.sync_local) {
          local_sync(warp < kEpiWarps && job.st_cw == 0, warp == kProducerWarp);
        } else {
          job_sync(warp < kEpiWarps && job.st_cw == 0, warp == kProducerWarp);
        }
        if (threadIdx.x == 0) trace_mark<TR>(tr, 13);
      }
    }
    // ---- K4: init (k == 0) or fused posterior + noise + inpaint + operand re-pack for this CTA's samples ----
    {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(lp.steps + k);
      uint32_t* dst = reinterpret_cast<uint32_t*>(&sstep);
      for (int i = threadIdx.x; i < (int)(sizeof(StepParams) / 4); i += kConvThreads) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    if (warp < kEpiWarps && s_hi > s_lo)
      ddpm_step_range<true>(sstep, s_lo * ht4, s_hi * ht4, threadIdx.x, kEpiThreads);
    if (lp.k4_local) local_sync(warp < kEpiWarps, warp == kProducerWarp);
    else job_sync(warp < kEpiWarps, warp == kProducerWarp);
  }
  if (threadIdx.x == 0) tma_store_wait_all();
  pair_end(warp, tmem_base, 512);
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device (context): done once for each device this process
// launches on, under a lock (host threads may drive different GPUs).
static cudaError_t ensure_smem_attr() {
  static std::mutex mu;
  static bool done[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(mu);
  if (dev >= 0 && dev < 64 && done[dev]) return cudaSuccess;
  const void* kernels[] = {(const void*)conv_gemm_kernel<false, false>, (const void*)conv_gemm_kernel<true, false>,
                           (const void*)conv_gemm_kernel<false, true>, (const void*)conv_gemm_kernel<true, true>,
                           (const void*)sample_loop_kernel<false, false, false, false>, (const void*)sample_loop_kernel<false, true, false, false>,
                           (const void*)sample_loop_kernel<true, false, false, false>, (const void*)sample_loop_kernel<true, true, false, false>,
                           (const void*)sample_loop_kernel<false, false, true, false>, (const void*)sample_loop_kernel<false, true, true, false>,
                           (const void*)sample_loop_kernel<true, false, true, false>, (const void*)sample_loop_kernel<true, true, true, false>,
                           (const void*)sample_loop_kernel<false, false, false, true>, (const void*)sample_loop_kernel<true, false, false, true>,
                           (const void*)sample_loop_kernel<false, true, false, true>, (const void*)sample_loop_kernel<true, true, false, true>,
                           (const void*)sample_loop_kernel<false, false, false, false, true>,
                           (const void*)sample_loop_kernel<true, false, false, false, true>,
                           (const void*)sample_loop_kernel<false, true, false, false, true>,
                           (const void*)sample_loop_kernel<true, true, false, false, true>};
  for (const void* k : kernels) {
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
    if (e != cudaSuccess) break;
  }
  if (e == cudaSuccess && dev >= 0 && dev < 64) done[dev] = true;
  return e;
}

cudaError_t launch_conv_job(const ConvMaps& maps, const ConvJob& job, int n_groups, int n_slices, size_t smem_bytes,
                            cudaStream_t stream) {
  cudaError_t e = ensure_smem_attr();
  if (e != cudaSuccess) return e;
  // one CTA pair (cluster of 2 along x) per 16-sample group; two pairs (cluster of 4) when the positions are split
  dim3 grid((job.ls ? 4 : 2) * n_groups, n_slices);
  const bool rd = job.n_rounds > 1 || job.st_wave;
  if (job.ls) {
    if (rd) conv_gemm_kernel<true, true><<<grid, kConvThreads, smem_bytes, stream>>>(maps, job);
    else conv_gemm_kernel<false, true><<<grid, kConvThreads, smem_bytes, stream>>>(maps, job);
  } else {
    if (rd) conv_gemm_kernel<true, false><<<grid, kConvThreads, smem_bytes, stream>>>(maps, job);
    else conv_gemm_kernel<false, false><<<grid, kConvThreads, smem_bytes, stream>>>(maps, job);
  }
  return cudaGetLastError();
}

cudaError_t launch_sample_loop(const LoopParams& lp, size_t smem_bytes, cudaStream_t stream) {
  cudaError_t e = ensure_smem_attr();
  if (e != cudaSuccess) return e;
  bool rounds = false;   // any layer that passes through TMEM in rounds (horizons above 256) selects the RD = true variant
  for (int i = 0; i < lp.n_jobs; ++i) rounds = rounds || lp.jobs[i].n_rounds > 1 || lp.jobs[i].st_wave;
  const bool ls = lp.n_jobs > 0 && lp.jobs[0].ls;   // the plan splits every layer or none
  const int grid = (ls ? 4 : 2) * lp.G;
  bool at = false;
  for (int i = 0; i < lp.n_jobs; ++i) at = at || lp.jobs[i].kind != JOB_CONV;
  if (at) {
    if (ls) return cudaErrorInvalidValue;   // (dm_plan_create never splits the positions of an attention model)
    if (lp.trace) { if (rounds) sample_loop_kernel<true, true, false, true><<<grid, kConvThreads, smem_bytes, stream>>>(lp);
                    else sample_loop_kernel<true, false, false, true><<<grid, kConvThreads, smem_bytes, stream>>>(lp); }
    else { if (rounds) sample_loop_kernel<false, true, false, true><<<grid, kConvThreads, smem_bytes, stream>>>(lp);
           else sample_loop_kernel<false, false, false, true><<<grid, kConvThreads, smem_bytes, stream>>>(lp); }
    return cudaGetLastError();
  }
  if (lp.n_jobs > 0 && lp.jobs[0].mc) {   // (the plan sets it on every conv job or on none; it implies !ls, !at, G even)
    if (ls || (lp.G & 1)) return cudaErrorInvalidValue;
#define DM_LOOP_MC(TR_, RD_) sample_loop_kernel<TR_, RD_, false, false, true><<<grid, kConvThreads, smem_bytes, stream>>>(lp)
    if (lp.trace) { if (rounds) DM_LOOP_MC(true, true); else DM_LOOP_MC(true, false); }
    else { if (rounds) DM_LOOP_MC(false, true); else DM_LOOP_MC(false, false); }
#undef DM_LOOP_MC
    return cudaGetLastError();
  }
#define DM_LOOP(TR_, RD_, LS_) sample_loop_kernel<TR_, RD_, LS_, false><<<grid, kConvThreads, smem_bytes, stream>>>(lp)
  if (ls) {
    if (lp.trace) { if (rounds) DM_LOOP(true, true, true); else DM_LOOP(true, false, true); }
    else { if (rounds) DM_LOOP(false, true, true); else DM_LOOP(false, false, true); }
  } else {
    if (lp.trace) { if (rounds) DM_LOOP(true, true, false); else DM_LOOP(true, false, false); }
    else { if (rounds) DM_LOOP(false, true, false); else DM_LOOP(false, false, false); }
  }
#undef DM_LOOP
  return cudaGetLastError();
}

// Clusters of two / four CTAs of the loop kernel the device can hold at once (GPC boundaries and disabled SMs decide, not just
// the SM count): the L-split and multicast layouts are only worth it when every cluster of the plan is resident in ONE wave.
cudaError_t max_resident_clusters(int cluster_size, size_t smem_bytes, int* n) {
  cudaError_t e = ensure_smem_attr();
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster_size * 148);
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem_bytes;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster_size; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (cluster_size == 4) return cudaOccupancyMaxActiveClusters(n, (const void*)sample_loop_kernel<false, false, false, false, true>, &cfg);
  return cudaOccupancyMaxActiveClusters(n, (const void*)sample_loop_kernel<false, false, false, false, false>, &cfg);
}

cudaError_t read_device_error(unsigned int* code) {
  return cudaMemcpyFromSymbol(code, g_dm_device_error, sizeof(unsigned int));
}

}  // namespace dm
