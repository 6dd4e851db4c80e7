// Descriptor of one Conv1d-family layer launch (K1 + K2 of SURVEY.md §2b), shared by host plan builder and
// the device kernel.  One launch computes, for every sample group, a conv (k=5 pad 2 | k=3 stride 2 |
// k=1 | ConvTranspose k=4 stride 2) as an implicit GEMM on tcgen05, followed in the same kernel by the
// epilogue of the reference block: bias, GroupNorm(8), Mish, + time bias, + residual.
//
// Reference semantics restated here:
//   Conv1dBlock            diffuser/models/helpers.py:50-69
//   ResidualTemporalBlock  diffuser/models/temporal.py:18-46
//   Downsample1d/Upsample1d diffuser/models/helpers.py:32-47
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace dm {

constexpr int kGroupSamples = 16;   // samples per activation group: layout [plane][G][L][16][C] fp16; one CTA PAIR per group
constexpr int kMaxTaps = 5;
constexpr int kEpiWarps = 16;       // epilogue warps 0-15: TMEM lane quarter = warp % 4, work slice = warp / 4
constexpr int kProducerWarp = 16;   // one elected lane issues TMA
constexpr int kMmaWarp = 17;        // one elected lane (leader CTA only) issues tcgen05.mma.cta_group::2; the warp owns the TMEM allocation
constexpr int kMmaWarpB = 18;       // second MMA issuer: one thread alone cannot keep the tensor pipe busy (see mma_unit)
constexpr int kConvThreads = (kEpiWarps + 3) * 32;

struct ConvTap {       // bytes: 80 of these descriptors travel as ONE kernel parameter block (32 KB limit)
  signed char wtap;    // index into the packed weight's tap dimension
  signed char shift;   // l offset of this tap's view inside the A box (in l positions)
  signed char phase;   // accumulator set (ConvTranspose: 0 = even outputs, 1 = odd outputs)
  signed char first;   // 1 if this is the first tap accumulated into its phase
};

enum { EPI_BIAS = 0, EPI_GN_MISH = 1 };
// ConvJob::kind.  The attention stages (attn_stages.cuh) are jobs of the same list and reuse the descriptor's activation
// fields; everything about tiles, taps and TMEM is meaningless for them.
enum { JOB_CONV = 0, JOB_ATTN_NORM = 1, JOB_ATTN_CORE = 2 };

// TMA descriptors of one layer; they live in global memory (or in a __grid_constant__ kernel parameter).
struct ConvMaps {
  CUtensorMap a_map[2];   // 5-D (C, 16, L, G, plane) fp16, one per input source (2 = channel concat)
  CUtensorMap w_map;      // 5-D (Cc, CoutPad, chunk, ntaps, plane) fp16: chunk-major, a tile is one contiguous block per plane
  CUtensorMap o_map;      // 5-D store map of the output activation (box = one staged plane x channel block)
  CUtensorMap r_map;      // 5-D load map of the residual activation, same box as o_map
  CUtensorMap x_map;      // 5-D weights of the fused auxiliary 1x1 conv (the block's residual_conv), layout and box as w_map
  CUtensorMap p_map;      // 5-D store map of the auxiliary output activation, same box as o_map
  // weight multicast (ConvJob::mc): boxes of a QUARTER of the tile's N rows, one plane each
  CUtensorMap wq_map;     // as w_map, box (Cc, Nc / 4, 1, 1, 1)
  CUtensorMap xq_map;     // as x_map, box (Cc, Nc / 4, 1, 1, 1)
};

// Plain-data layer descriptor.  It is always read from kernel-parameter (constant-bank) space, so every field
// is warp-uniform for the compiler: the MMA / TMA issue loops stay on the uniform datapath.
struct ConvJob {
  int nsrc;
  int src_chunks[2];      // channel chunks per source
  int Cc;                 // channels per chunk (64 -> SWIZZLE_128B rows, 32 -> SWIZZLE_64B rows)
  int row_bytes;          // Cc * 2
  int ksteps;             // k-steps (16 channels) issued per chunk: Cc / 16, less when the only chunk is zero padded
  int swz;                // UMMA layout code (2 = 128B, 4 = 64B)
  int S;                  // samples per CTA: always 8 (CTA r of the pair owns samples 8r .. 8r+7 of the group)
  int mrows;              // GEMM rows per CTA per MMA: 128 (tcgen05.mma.cta_group::2 with M = 256) or 64 (M = 128, L <= 8 layers)
  int ncols;              // TMEM columns of one M tile of one accumulator set: Nc (mrows 128) or Nc / 2 (mrows 64: the
                          // pair MMA puts channels [0, Nc/2) on lanes 0-63 and [Nc/2, Nc) on lanes 64-127)
  int l_stride;           // 1, or 2 for the strided down-sampling conv
  int halo_lo;            // A box starts at l = l_stride * j0 - halo_lo
  int box_l;              // l positions in the A box
  int Lt;                 // M positions (j) per l-tile; Lt * S is a multiple of mrows
  int n_ltiles;
  int Mt;                 // MMA tiles (mrows rows per CTA) per l-tile
  int Lj;                 // number of valid M positions (Lout, or Lin for ConvTranspose)
  int Lout;               // output length (of this CTA pair's share, see `ls`)
  // L-split (ls = 1; small batches at long horizons): a cluster of FOUR CTAs = two pairs owns the group, pair h works on
  // positions [h * L / 2, (h + 1) * L / 2) of every layer.  Lj / Lout / tiles describe ONE half; the kernel adds the offsets:
  // (flag `ls` in the byte block at the end)
  int ls_in;              // input positions between the halves' A boxes  (l_stride * Lj)
  int ls_out;             // output positions between the halves' rows    (Lout)
  int Lfull_out;          // length of the whole output activation (row addressing); == Lout when ls == 0
  int gn_L;               // positions one GroupNorm statistic covers (both halves); == Lj when ls == 0
  int ntaps;
  ConvTap taps[kMaxTaps];
  int nphase;
  int Nc;                 // output channels per unit == MMA N; each CTA of the pair stages Nc / 2 rows of every weight tile
  int n_a_stages, n_w_stages;
  int a_stage_bytes, w_stage_bytes, a_plane_bytes, w_plane_bytes;
  int a_tx_bytes, w_tx_bytes;   // bytes ONE CTA's loads deliver per stage (the leader's barrier expects twice that)
  // TMEM rounds.  The whole per-sample conv output normally stays in TMEM until the GroupNorm statistics are known; when
  // its tiles exceed the 512 columns (horizons above 256) the l-tiles are processed lt_round at a time: the epilogue drains
  // each round (statistics + raw fp32 values to `raw`), the MMAs reuse the columns, and pass 2 reads `raw` back.
  int lt_round;           // l-tiles per round (== n_ltiles when everything fits)
  int n_rounds;           // ceil(n_ltiles / lt_round)
  float* raw;             // [G][Lj][16][raw_C] fp32 scratch of the pre-normalisation conv output (rounds only)
  int raw_C;
  int tmem_cols;
  int mma_mode;           // 0: warp A issues every MMA; 1: the two issuer warps split the k-steps of every tap, each
                          // with its own accumulator set (see mma_unit)
  int acc2_off;
  // epilogue
  int epi;
  int Cout;               // real output channels
  int cg;                 // channels per GroupNorm group
  int B;                  // real batch (masks the fp32 output)
  const float* bias;      // [CoutPad]
  const float* gamma;     // [CoutPad]
  const float* beta;      // [CoutPad]
  const float* tbias;     // time bias rows, added after Mish (temporal.py:44); null if none
  int tbias_stride;       // 0: one row shared by all samples (uniform t); else per-sample row stride
  const __half* res;      // residual planes in activation layout, added last (temporal.py:46); null if none
  int res_C;
  long long res_plane_stride;
  __half* out;            // output planes in activation layout; null if none
  int out_C;
  long long out_plane_stride;
  float* out_f32;         // public layout [B][Lout][out_f32_C]; null if none
  int out_f32_C;
  int f32_stride;         // floats per sample in the smem staging of the fp32 output (0: direct global stores)
  // staged epilogue: the unit's output tile is assembled in shared memory (same swizzle as a TMA box) and
  // written with TMA stores; the residual tile is TMA-loaded into the same buffer.  st_cw == 0: direct path.
  int st_cw;              // channels per staged block (64 -> SWIZZLE_128B rows, 32 -> SWIZZLE_64B rows)
  int st_nq;              // channel blocks per unit (Nc / st_cw)
  int st_rows;            // rows (l, b) per block = nphase * n_ltiles * Mt * 128
  int st_lbox;            // l extent of the store box = st_rows / S
  int st_blk_stride;      // bytes between staged blocks: kStageGap zero rows | st_rows data rows | kStageGap zero rows
  // resident links (fused loop only): when the next layer's only input is this CTA's staged output tile, the
  // tile stays in shared memory (zero halo gaps around each block) and is read in place by the next layer's
  // MMAs: no TMA round trip through L2, no cluster barrier, and the TMA store of the tile is not waited for.
  int a_res_off;          // byte offset (from the smem base) of chunk 0 / plane 0 of the A view (row l = -halo_lo)
  int a_res_chunk;        // bytes between channel chunks of the resident tile
  int a_res_plane;        // bytes between the hi and lo planes of the resident tile
  int w_off_res;          // weight ring offset when a_res (beyond every staged tile)
  int pre_next;           // weight loads of the NEXT layer's first chunk the producer may issue during this epilogue
  // auxiliary 1x1 conv sharing this layer's A operand (ResidualTemporalBlock.residual_conv, temporal.py:34-35,46):
  // one extra weight tile per channel chunk, accumulated into its own TMEM columns, bias-only epilogue
  int aux_shift;          // l offset of the 1x1 tap inside the A box (= halo_lo)
  int aux_col0;           // first TMEM column of the auxiliary accumulators
  const float* aux_bias;  // [CoutPad]
  // one-byte flags, kept together (80 descriptors share the 32 KB kernel-parameter block)
  unsigned char kind;               // JOB_CONV, or one of the attention stages
  unsigned char mc;                 // fused loop only: clusters of four = TWO groups; the pairs share every weight tile by TMA multicast
  unsigned char ls;                 // 1: L-split job (see ls_in / ls_out above)
  unsigned char cg_pow2;            // cg is a power of two (lets the epilogue fold partial sums by shifts)
  unsigned char a_res;              // this layer's A operand is the previous layer's staged tile
  unsigned char sync_local;         // the next layer only reads what THIS CTA wrote (both layers S = 8): block barrier, no cluster barrier
  unsigned char next_res;           // the next layer reads this layer's staged tile in place (zero the gaps, defer the store wait)
  unsigned char gaps_clean;         // resident producer whose staging area was last written by an identically laid-out tile (the previous
                          // layer's, which this layer read in place): the zero halo gaps are still zero, skip re-zeroing them
  unsigned char no_store;           // fused loop only: the resident consumer is the ONLY reader of this layer's output, so the tile
                          // never leaves shared memory (no TMA store to the global activation planes)
  unsigned char aux;                // 1: the unit carries the auxiliary 1x1 conv (aux_* above)
  unsigned char st_wave;            // 1: the tile exceeds the staging budget as a whole; it is staged and stored in waves of four
                          // (tile, 32-channel) units through two buffers of 2 * st_nq * st_blk_stride bytes (st_rows = rows per wave)
  unsigned char acc2;               // 1: a second accumulator set at TMEM column acc2_off is added by the epilogue
  // fused-loop bookkeeping
  int slices;             // number of N slices (units per sample sub-group)
  int tb_off;             // offset of the layer's time bias inside a time-bias table row, or -1
};

constexpr int kStageGap = 2048;    // bytes of zero rows before and after every staged block (conv halo)
constexpr int kMaxLoopJobs = 80;   // layers of one U-Net inside the fused-loop kernel parameter block

}  // namespace dm
