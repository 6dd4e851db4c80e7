// Plan builder + C ABI for the U-Net / sampling-loop part of the hot path.  Host C++ only; the arithmetic is
// in conv_kernel.cu and step_kernels.cu.
#include "plan.h"

#include <math.h>
#include <algorithm>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/dm_b200.h"
#include "attn_stages.cuh"

namespace dm {

cudaError_t launch_conv_job(const ConvMaps& maps, const ConvJob& job, int n_groups, int n_slices, size_t smem_bytes,
                            cudaStream_t stream);
cudaError_t read_device_error(unsigned int* code);
cudaError_t max_resident_clusters(int cluster_size, size_t smem_bytes, int* n);

static thread_local char g_err[512] = "";
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
const char* last_error() { return g_err; }

#define DM_CUDA(expr)                                                                              \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) return fail(DM_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

int check_device() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(DM_ERR_DEVICE, "no CUDA device: this library has no CPU fallback");
  }
  int dev = 0, major = 0;
  DM_CUDA(cudaGetDevice(&dev));
  DM_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) return fail(DM_ERR_DEVICE, "device compute capability %d.x is not sm_100 (B200)", major);
  return DM_OK;
}

// Every entry point that takes a plan runs on the plan's device: a different current device would launch kernels on
// a GPU that holds none of the plan's memory.
int check_plan_device(const dm_plan* P) {
  int dev = -1;
  DM_CUDA(cudaGetDevice(&dev));
  if (dev != P->device)
    return fail(DM_ERR_STATE, "plan was created on device %d but the current device is %d (use cudaSetDevice / "
                              "torch.cuda.device)", P->device, dev);
  return DM_OK;
}

// ---- TMA descriptor encoding through the driver entry point (no link-time libcuda dependency) ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}
int encode_map(CUtensorMap* m, void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swz) {
  EncodeTiledFn fn = get_encode();
  if (!fn) return fail(DM_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUtensorMapSwizzle sw = swz == 2 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, rank, base, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(DM_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return DM_OK;
}

typedef CUresult (*ReplaceAddressFn)(CUtensorMap*, void*);
static ReplaceAddressFn get_replace_address() {
  static ReplaceAddressFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapReplaceAddress", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<ReplaceAddressFn>(p);
  }
  return fn;
}

// ---- allocation helpers ----
static int dev_alloc(dm_plan* P, void** p, size_t bytes, bool zero = true) {
  DM_CUDA(cudaMalloc(p, bytes));
  P->allocs.push_back(*p);
  if (zero) DM_CUDA(cudaMemset(*p, 0, bytes));
  return DM_OK;
}
static int upload(dm_plan* P, const std::vector<float>& h, float** d) {
  int rc = dev_alloc(P, reinterpret_cast<void**>(d), h.size() * sizeof(float), false);
  if (rc) return rc;
  DM_CUDA(cudaMemcpy(*d, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
  return DM_OK;
}
static int new_act(dm_plan* P, int C, int L, int* idx) {
  Act a;
  a.C = C;
  a.L = L;
  a.plane = (long long)P->G * L * kGroupSamples * C;
  int rc = dev_alloc(P, reinterpret_cast<void**>(&a.p), (size_t)a.plane * 2 * sizeof(__half));
  if (rc) return rc;
  P->acts.push_back(a);
  *idx = (int)P->acts.size() - 1;
  return DM_OK;
}

typedef std::map<std::string, HostTensor> WeightMap;
static int find_w(const WeightMap& W, const std::string& name, std::vector<int64_t> shape, const float** out) {
  auto it = W.find(name);
  if (it == W.end()) return fail(DM_ERR_WEIGHT, "state_dict tensor '%s' missing", name.c_str());
  if (it->second.shape != shape) return fail(DM_ERR_WEIGHT, "state_dict tensor '%s' has the wrong shape", name.c_str());
  *out = it->second.data;
  return DM_OK;
}

// Weight loads of `nx`'s first channel chunk that may be issued while `cur`'s epilogue runs: ring slots of nx
// that lie beyond cur's staged output tile (cur's operand stages are dead by then).
static int prefetchable_w_loads(const ConvJob& cur, const ConvJob& nx) {
  int cur_st = cur.st_cw ? (cur.st_wave ? 2 : 1) * 2 * cur.st_nq * cur.st_blk_stride : 0;
  if (cur.f32_stride && cur.S * cur.f32_stride * 4 > cur_st) cur_st = cur.S * cur.f32_stride * 4;
  const int w_off = nx.a_res ? nx.w_off_res : nx.n_a_stages * nx.a_stage_bytes;
  const int lim = nx.ntaps < nx.n_w_stages ? nx.ntaps : nx.n_w_stages;
  return w_off >= cur_st ? lim : 0;   // slots 0..lim-1 start at w_off: all of them or none
}

enum ConvKind { CONV_K5 = 0, CONV_DOWN = 1, CONV_K1 = 2, CONV_T4 = 3 };

struct ConvSpec {
  std::string name;      // debug name
  std::string wkey;      // state_dict prefix of the conv: "<wkey>.weight", "<wkey>.bias"
  std::string gnkey;     // GroupNorm prefix or ""
  ConvKind kind;
  int src[2] = {-1, -1}; // act indices
  int src_real_c[2] = {0, 0};  // real (unpadded) channels each source contributes to the weight's Cin
  int nsrc = 1;
  int Cout = 0;
  int tb_off = -1;
  int res = -1;          // residual act index
  bool want_act_out = true;
  bool f32_out = false;  // write plan->model_out style fp32 [B,L,Cout]
  float* f32_dst = nullptr;   // with f32_out: the destination when it is not plan->model_out (attention: the q|k|v rows)
  std::string aux_wkey;  // fused auxiliary 1x1 conv over the same input ("<key>.weight/.bias"), or ""
  bool no_bias = false;  // Conv1d(bias=False): LinearAttention.to_qkv (helpers.py:114)
  bool hidden = false;   // intermediate of an attention block (no entry of its own in the oracle's per-layer record)
};

static int round_up(int a, int b) { return (a + b - 1) / b * b; }

// Adds one conv launch to the plan; returns the output act index in *out_idx (or -1).
static int add_conv(dm_plan* P, const WeightMap& W, const ConvSpec& sp, int* out_idx, int* aux_idx = nullptr) {
  Layer ly;
  ConvJob& j = ly.job;
  memset(&j, 0, sizeof(j));
  memset(&ly.maps, 0, sizeof(ly.maps));
  ly.name = sp.name;
  ly.hidden = sp.hidden;
  ly.tb_off = sp.tb_off;
  for (int s = 0; s < sp.nsrc; ++s) ly.src_acts[s] = sp.src[s];
  ly.res_act = sp.res;

  const Act& a0 = P->acts[sp.src[0]];
  const int Lin = a0.L;
  int ktaps = sp.kind == CONV_K5 ? 5 : sp.kind == CONV_DOWN ? 3 : sp.kind == CONV_K1 ? 1 : 4;
  const int Lout = sp.kind == CONV_DOWN ? Lin / 2 : sp.kind == CONV_T4 ? Lin * 2 : Lin;
  // L-split: all tile geometry below describes ONE pair's half of the positions (Lj, Lo); activations keep their full length
  const int ls = P->lsplit ? 1 : 0;
  const int Lj_full = sp.kind == CONV_T4 ? Lin : Lout;
  if (ls && (Lj_full % 2 || Lout % 2)) return fail(DM_ERR_STATE, "%s: odd length, cannot split the positions", sp.name.c_str());
  const int Lj = Lj_full >> ls;
  const int Lo = Lout >> ls;

  // channel chunking
  int cin_tot = 0, cin_real = 0;
  bool all64 = true;
  for (int s = 0; s < sp.nsrc; ++s) {
    const Act& a = P->acts[sp.src[s]];
    if (a.L != Lin) return fail(DM_ERR_ARG, "%s: concat sources differ in length", sp.name.c_str());
    if (a.C % 64) all64 = false;
    cin_tot += a.C;
    cin_real += sp.src_real_c[s];
  }
  j.Cc = all64 ? 64 : 32;
  j.row_bytes = j.Cc * 2;
  j.swz = all64 ? 2 : 4;
  j.nsrc = sp.nsrc;
  for (int s = 0; s < sp.nsrc; ++s) j.src_chunks[s] = P->acts[sp.src[s]].C / j.Cc;
  j.ksteps = j.Cc / 16;
  if (sp.nsrc == 1 && j.src_chunks[0] == 1 && sp.src_real_c[0] < j.Cc)   // e.g. the 35 qpos channels padded to 64
    j.ksteps = (sp.src_real_c[0] + 15) / 16;

  // tile shape.  A CTA always owns 8 samples of its group (GEMM row = l * 8 + b); the two CTAs of a cluster form a CTA
  // pair whose MMAs (tcgen05.mma.cta_group::2) span both CTAs' rows: M = 256 (128 rows per CTA), or M = 128 (64 rows per
  // CTA) for the layers with L <= 8.  Each CTA stages half of every weight tile (N / 2 rows); the pair shares them.
  const bool gn = !sp.gnkey.empty();
  const int cout_pad = round_up(sp.Cout, 32);
  const int cg = gn ? sp.Cout / 8 : 32;
  if (gn && (sp.Cout % 32 != 0)) return fail(DM_ERR_ARG, "%s: GroupNorm layers need Cout %% 32 == 0", sp.name.c_str());
  j.l_stride = sp.kind == CONV_DOWN ? 2 : 1;
  j.S = 8;
  j.mrows = (Lj * 8 <= 64 && cout_pad % 64 == 0) ? 64 : 128;
  const int pos_per_mt = j.mrows / j.S;
  int maxshift;
  if (sp.kind == CONV_K5) { j.halo_lo = 2; maxshift = 4; }
  else if (sp.kind == CONV_DOWN) { j.halo_lo = 1; maxshift = 2; }
  else if (sp.kind == CONV_K1) { j.halo_lo = 0; maxshift = 0; }
  else { j.halo_lo = 1; maxshift = 2; }
  const int mt_total = (Lj + pos_per_mt - 1) / pos_per_mt;
  auto box_for = [&](int mt) { return round_up(j.l_stride * (mt * pos_per_mt - 1) + maxshift + 1, 2); };
  auto stage_for = [&](int mt) { return 2 * box_for(mt) * j.S * j.row_bytes; };
  int Mt = mt_total;
  const int a_budget = 80 * 1024;
  while (Mt > 1 && (stage_for(Mt) > a_budget || box_for(Mt) > 256)) --Mt;
  j.n_ltiles = (mt_total + Mt - 1) / Mt;
  Mt = (mt_total + j.n_ltiles - 1) / j.n_ltiles;   // even split: no padding tiles beyond mt_total (TMEM budget)
  j.Mt = Mt;
  j.Lt = Mt * pos_per_mt;
  j.box_l = box_for(Mt);
  j.Lj = Lj;
  j.Lout = Lo;
  j.ls = ls;
  j.ls_in = ls ? j.l_stride * Lj : 0;
  j.ls_out = ls ? Lo : 0;
  j.Lfull_out = Lout;
  j.gn_L = Lj_full;
  // a half that is not a whole number of MMA tiles (H = 480: 30 positions at the deepest level): the padded rows of its
  // last tile must not be stored (they would land in the other half), so the store boxes below cover the valid rows only
  const bool ls_padded = ls && (Lj % pos_per_mt || j.n_ltiles * Mt * pos_per_mt != Lj);
  ly.ls_padded = ls_padded;
  j.a_plane_bytes = j.box_l * j.S * j.row_bytes;
  j.a_tx_bytes = 2 * j.a_plane_bytes;
  j.a_stage_bytes = round_up(j.a_tx_bytes, 1024);

  // taps
  j.nphase = 1;
  if (sp.kind == CONV_T4) {
    // ConvTranspose1d(k=4,s=2,p=1): out[2j] = W1 x[j] + W3 x[j-1]; out[2j+1] = W2 x[j] + W0 x[j+1]
    j.nphase = 2;
    j.ntaps = 4;
    j.taps[0] = {1, 1, 0, 1};
    j.taps[1] = {3, 0, 0, 0};
    j.taps[2] = {2, 1, 1, 1};
    j.taps[3] = {0, 2, 1, 0};
  } else {
    j.ntaps = ktaps;
    for (int t = 0; t < ktaps; ++t) j.taps[t] = {(signed char)t, (signed char)t, 0, (signed char)(t == 0 ? 1 : 0)};
  }

  // N split and TMEM budget.  N = 256 for the 64-row layers (M = 128 pair MMA: the widest N reads A least often), else 128.
  const int n_step = j.mrows == 64 ? 64 : 32;                  // mrows 64: each lane half holds Nc / 2 columns (32-column units)
  // (the attention block's 384-channel to_qkv layer uses N = 192 when it is a single M tile: two units instead of three,
  // and the output tile still fits the staged epilogue)
  const int n_cap = (j.mrows == 64 || (sp.hidden && mt_total == 1)) ? 256 : 128;
  int Nc = cout_pad < n_cap ? cout_pad : n_cap;
  while (cout_pad % Nc || (gn && Nc % cg)) Nc -= n_step;
  const int mt_all = j.n_ltiles * j.Mt;
  auto tile_cols = [&](int n) { return j.mrows == 64 ? n / 2 : n; };
  int acc_sets = j.nphase + (sp.aux_wkey.empty() ? 0 : 1);   // the auxiliary conv has its own accumulators
  if (!sp.aux_wkey.empty() && acc_sets * mt_all * tile_cols(Nc) > 512)
    return fail(DM_ERR_STATE, "%s: no TMEM room for the fused residual conv", sp.name.c_str());   // caller falls back
  while (j.nphase * mt_all * tile_cols(Nc) > 512 && Nc > n_step && (Nc / 2) % n_step == 0 && (!gn || (Nc / 2) % cg == 0)) Nc /= 2;
  if (Nc < 16 || (gn && Nc % cg)) return fail(DM_ERR_ARG, "%s: N slice does not hold whole GroupNorm groups", sp.name.c_str());
  j.Nc = Nc;
  j.ncols = tile_cols(Nc);
  // TMEM rounds (horizons above 256): when even the narrowest N slice does not fit, the l-tiles go through TMEM
  // lt_round at a time and the GroupNorm epilogue keeps the raw conv output in a global scratch between its two passes
  j.lt_round = j.n_ltiles;
  if (j.nphase * mt_all * j.ncols > 512) {
    if (ls) return fail(DM_ERR_STATE, "%s: TMEM rounds are not combined with the position split", sp.name.c_str());
    j.lt_round = 512 / (j.nphase * j.Mt * j.ncols);
    if (j.lt_round < 1)
      return fail(DM_ERR_ARG, "%s: one l-tile does not fit the 512 TMEM columns", sp.name.c_str());
    if (!sp.aux_wkey.empty())
      return fail(DM_ERR_STATE, "%s: no TMEM room for the fused residual conv", sp.name.c_str());   // caller falls back
  }
  j.n_rounds = (j.n_ltiles + j.lt_round - 1) / j.lt_round;
  const int mt_acc = j.lt_round * j.Mt;                        // M tiles per phase resident in TMEM at a time
  // two MMA issuer warps with one accumulator set each (conv_kernel.cu, mma_unit) when TMEM has room for it
  j.acc2 = 0;
  j.acc2_off = 0;
#ifdef DM_SINGLE_ISSUER
  if (false) {
#else
  if (j.ksteps >= 2 && (acc_sets + j.nphase) * mt_acc * j.ncols <= 512) {
#endif
    j.acc2 = 1;
    j.acc2_off = j.nphase * mt_acc * j.ncols;
    acc_sets += j.nphase;
  }
  ly.slices = cout_pad / Nc;
  j.slices = ly.slices;
  j.tb_off = sp.tb_off;
  int cols = 32;
  while (cols < acc_sets * mt_acc * j.ncols) cols *= 2;
  j.tmem_cols = cols;
  j.w_plane_bytes = (Nc / 2) * j.row_bytes;                     // this CTA's half of the weight tile
  j.w_tx_bytes = 2 * j.w_plane_bytes;
  j.w_stage_bytes = round_up(j.w_tx_bytes, 1024);

  // pipeline depth under a 200 KB shared-memory budget
  const int total_chunks = (j.src_chunks[0] + (sp.nsrc > 1 ? j.src_chunks[1] : 0)) * j.n_ltiles;
  j.n_a_stages = total_chunks > 1 ? 2 : 1;
  const int budget = 200 * 1024;
  if (j.n_a_stages * j.a_stage_bytes + j.w_stage_bytes > budget) j.n_a_stages = 1;
  int nw = (budget - j.n_a_stages * j.a_stage_bytes) / j.w_stage_bytes;
  const int w_loads = total_chunks * (j.ntaps + (sp.aux_wkey.empty() ? 0 : 1));
  if (nw > w_loads) nw = w_loads;
  if (nw > 8) nw = 8;
  if (nw < 1) return fail(DM_ERR_ARG, "%s: tile does not fit in shared memory", sp.name.c_str());
  j.n_w_stages = nw;
  ly.smem = (size_t)j.n_a_stages * j.a_stage_bytes + (size_t)nw * j.w_stage_bytes + 1024;

  // ---- pack weights: planes [2][ktaps][chunk][cout_pad][Cc] fp16, zero padded.  Chunk-major: the Nc x Cc tile one
  //      TMA load fetches is ONE contiguous block per plane (rows 128 / 64 bytes apart), not Nc pieces strided by the
  //      layer's full input width ----
  const float* wsrc = nullptr;
  int rc;
  if (sp.kind == CONV_T4) rc = find_w(W, sp.wkey + ".weight", {cin_real, sp.Cout, ktaps}, &wsrc);
  else rc = find_w(W, sp.wkey + ".weight", {sp.Cout, cin_real, ktaps}, &wsrc);
  if (rc) return rc;
  const float* bsrc = nullptr;
  if (!sp.no_bias && (rc = find_w(W, sp.wkey + ".bias", {sp.Cout}, &bsrc))) return rc;
  const int n_chunks_tot = cin_tot / j.Cc;
  const size_t wplane = (size_t)ktaps * cout_pad * cin_tot;
  auto w_index = [&](int k, int co, int col) {   // col: channel position inside the concatenated, padded input
    return (((size_t)k * n_chunks_tot + col / j.Cc) * cout_pad + co) * j.Cc + col % j.Cc;
  };
  std::vector<__half> wp(2 * wplane, __float2half_rn(0.f));
  {
    int col0 = 0, real0 = 0;
    for (int s = 0; s < sp.nsrc; ++s) {
      for (int ci = 0; ci < sp.src_real_c[s]; ++ci) {
        for (int co = 0; co < sp.Cout; ++co) {
          for (int k = 0; k < ktaps; ++k) {
            const int cr = real0 + ci;
            const float v = sp.kind == CONV_T4 ? wsrc[((size_t)cr * sp.Cout + co) * ktaps + k]
                                               : wsrc[((size_t)co * cin_real + cr) * ktaps + k];
            const __half hi = __float2half_rn(v);
            const __half lo = __float2half_rn(v - __half2float(hi));
            const size_t o = w_index(k, co, col0 + ci);
            wp[o] = hi;
            wp[wplane + o] = lo;
          }
        }
      }
      col0 += P->acts[sp.src[s]].C;
      real0 += sp.src_real_c[s];
    }
  }
  __half* wdev = nullptr;
  if ((rc = dev_alloc(P, reinterpret_cast<void**>(&wdev), wp.size() * sizeof(__half), false))) return rc;
  DM_CUDA(cudaMemcpy(wdev, wp.data(), wp.size() * sizeof(__half), cudaMemcpyHostToDevice));
  {
    const uint64_t dims[5] = {(uint64_t)j.Cc, (uint64_t)cout_pad, (uint64_t)n_chunks_tot, (uint64_t)ktaps, 2};
    const uint64_t str[4] = {(uint64_t)j.Cc * 2, (uint64_t)cout_pad * j.Cc * 2, (uint64_t)cin_tot * cout_pad * 2,
                             (uint64_t)wplane * 2};
    const uint32_t box[5] = {(uint32_t)j.Cc, (uint32_t)(Nc / 2), 1, 1, 2};   // per CTA: half of the tile's N rows
    if ((rc = encode_map(&ly.maps.w_map, wdev, 5, dims, str, box, j.swz))) return rc;
    if (Nc % 32 == 0) {   // weight multicast: a quarter of the rows (whole swizzle atoms), one plane per box
      const uint32_t qbox[5] = {(uint32_t)j.Cc, (uint32_t)(Nc / 4), 1, 1, 1};
      if ((rc = encode_map(&ly.maps.wq_map, wdev, 5, dims, str, qbox, j.swz))) return rc;
      ly.mc_ok = true;
    }
  }
  for (int s = 0; s < sp.nsrc; ++s) {
    const Act& a = P->acts[sp.src[s]];
    const uint64_t dims[5] = {(uint64_t)a.C, kGroupSamples, (uint64_t)a.L, (uint64_t)P->G, 2};
    const uint64_t str[4] = {(uint64_t)a.C * 2, (uint64_t)a.C * 2 * kGroupSamples,
                             (uint64_t)a.C * 2 * kGroupSamples * a.L, (uint64_t)a.plane * 2};
    const uint32_t box[5] = {(uint32_t)j.Cc, (uint32_t)j.S, (uint32_t)j.box_l, 1, 2};
    if ((rc = encode_map(&ly.maps.a_map[s], a.p, 5, dims, str, box, j.swz))) return rc;
  }

  // ---- epilogue parameters ----
  std::vector<float> hb(cout_pad, 0.f), hg(cout_pad, 1.f), hbe(cout_pad, 0.f);
  for (int c = 0; c < sp.Cout && bsrc; ++c) hb[c] = bsrc[c];
  if (gn) {
    const float *gs = nullptr, *bs = nullptr;
    if ((rc = find_w(W, sp.gnkey + ".weight", {sp.Cout}, &gs))) return rc;
    if ((rc = find_w(W, sp.gnkey + ".bias", {sp.Cout}, &bs))) return rc;
    for (int c = 0; c < sp.Cout; ++c) { hg[c] = gs[c]; hbe[c] = bs[c]; }
  }
  float *db, *dg, *dbe;
  if ((rc = upload(P, hb, &db)) || (rc = upload(P, hg, &dg)) || (rc = upload(P, hbe, &dbe))) return rc;
  j.epi = gn ? EPI_GN_MISH : EPI_BIAS;
  j.Cout = sp.Cout;
  j.cg = cg;
  j.cg_pow2 = (cg & (cg - 1)) == 0;
  j.B = P->B;
  j.bias = db;
  j.gamma = dg;
  j.beta = dbe;
  j.tbias = nullptr;
  j.tbias_stride = 0;
  if (sp.res >= 0) {
    const Act& r = P->acts[sp.res];
    if (r.L != Lout || r.C < cout_pad) return fail(DM_ERR_ARG, "%s: residual shape mismatch", sp.name.c_str());
    j.res = r.p;
    j.res_C = r.C;
    j.res_plane_stride = r.plane;
  }
  *out_idx = -1;
  if (sp.want_act_out) {
    int oi;
    if ((rc = new_act(P, cout_pad, Lout, &oi))) return rc;
    const Act& o = P->acts[oi];
    j.out = o.p;
    j.out_C = o.C;
    j.out_plane_stride = o.plane;
    ly.out_act = oi;
    *out_idx = oi;
  }
  if (sp.f32_out) {
    j.out_f32 = sp.f32_dst ? sp.f32_dst : P->model_out;
    j.out_f32_C = sp.Cout;
    const int per_sample = Lo * sp.Cout;
    if (per_sample % 4 == 0 && ly.slices == 1 && j.n_rounds == 1 && (size_t)j.S * (per_sample + 4) * 4 <= (size_t)budget) {
      j.f32_stride = per_sample + 4;
      const size_t need = (size_t)j.S * j.f32_stride * 4 + 1024;
      if (ly.smem < need) ly.smem = need;
    }
  }
  // ---- staged epilogue (TMA store of the output tile, TMA load of the residual tile) ----
  j.st_cw = 0;
  size_t st_bytes = 0;
  if (sp.want_act_out) {
    const int rows = j.nphase * mt_all * j.mrows;
    const int lbox = ls ? Lo : rows / j.S;           // (unsplit: TMA clips a padded box at the end of the activation)
    const int cw = Nc % 64 == 0 ? 64 : 32;
    const int blk_stride = rows * cw * 2 + 2 * kStageGap;
    st_bytes = (size_t)2 * (Nc / cw) * blk_stride;
    if (lbox <= 256 && st_bytes <= (size_t)budget && j.n_rounds == 1) {
      j.st_cw = cw;
      j.st_nq = Nc / cw;
      j.st_rows = lbox * j.S;
      j.st_lbox = lbox;
      j.st_blk_stride = blk_stride;
      if (ly.smem < st_bytes + 1024) ly.smem = st_bytes + 1024;
      const int st_swz = j.st_cw == 64 ? 2 : 4;
      const uint32_t box[5] = {(uint32_t)j.st_cw, (uint32_t)j.S, (uint32_t)lbox, 1, 1};
      {
        const Act& o = P->acts[ly.out_act];
        const uint64_t dims[5] = {(uint64_t)o.C, kGroupSamples, (uint64_t)o.L, (uint64_t)P->G, 2};
        const uint64_t str[4] = {(uint64_t)o.C * 2, (uint64_t)o.C * 2 * kGroupSamples,
                                 (uint64_t)o.C * 2 * kGroupSamples * o.L, (uint64_t)o.plane * 2};
        if ((rc = encode_map(&ly.maps.o_map, o.p, 5, dims, str, box, st_swz))) return rc;
      }
      if (sp.res >= 0) {
        const Act& r = P->acts[sp.res];
        const uint64_t dims[5] = {(uint64_t)r.C, kGroupSamples, (uint64_t)r.L, (uint64_t)P->G, 2};
        const uint64_t str[4] = {(uint64_t)r.C * 2, (uint64_t)r.C * 2 * kGroupSamples,
                                 (uint64_t)r.C * 2 * kGroupSamples * r.L, (uint64_t)r.plane * 2};
        if ((rc = encode_map(&ly.maps.r_map, r.p, 5, dims, str, box, st_swz))) return rc;
      }
    } else if (j.n_rounds == 1 && j.nphase == 1 && j.mrows == 128 && (Nc == 32 || Nc == 64 || Nc == 128) &&
               !(ls && (ls_padded || (mt_all * (Nc / 32)) % 4))) {   // (split: whole waves only, else the direct stores below)
      // big tile: wave-staged epilogue (conv_kernel.cu, pass 2).  One wave = 4 units of 128 rows x 32 channels.
      const int nchunks = Nc / 32;
      const int rows_w = 512 / nchunks;
      j.st_wave = 1;
      j.st_cw = cw;
      j.st_nq = Nc / cw;
      j.st_rows = rows_w;
      j.st_lbox = rows_w / j.S;
      j.st_blk_stride = rows_w * cw * 2;
      st_bytes = (size_t)2 * 2 * j.st_nq * j.st_blk_stride;        // two wave buffers
      if (ly.smem < st_bytes + 1024) ly.smem = st_bytes + 1024;
      const int st_swz = cw == 64 ? 2 : 4;
      const uint32_t box[5] = {(uint32_t)cw, (uint32_t)j.S, (uint32_t)j.st_lbox, 1, 1};
      {
        const Act& o = P->acts[ly.out_act];
        const uint64_t dims[5] = {(uint64_t)o.C, kGroupSamples, (uint64_t)o.L, (uint64_t)P->G, 2};
        const uint64_t str[4] = {(uint64_t)o.C * 2, (uint64_t)o.C * 2 * kGroupSamples,
                                 (uint64_t)o.C * 2 * kGroupSamples * o.L, (uint64_t)o.plane * 2};
        if ((rc = encode_map(&ly.maps.o_map, o.p, 5, dims, str, box, st_swz))) return rc;
      }
      if (sp.res >= 0) {
        const Act& r = P->acts[sp.res];
        const uint64_t dims[5] = {(uint64_t)r.C, kGroupSamples, (uint64_t)r.L, (uint64_t)P->G, 2};
        const uint64_t str[4] = {(uint64_t)r.C * 2, (uint64_t)r.C * 2 * kGroupSamples,
                                 (uint64_t)r.C * 2 * kGroupSamples * r.L, (uint64_t)r.plane * 2};
        if ((rc = encode_map(&ly.maps.r_map, r.p, 5, dims, str, box, st_swz))) return rc;
      }
    } else {
      st_bytes = 0;
    }
  }
  // ---- fused auxiliary 1x1 conv (ResidualTemporalBlock.residual_conv): weights, bias, output act + store map ----
  if (!sp.aux_wkey.empty()) {
    if (!gn || sp.kind != CONV_K5 || sp.res >= 0 || j.st_cw == 0 || j.st_wave)
      return fail(DM_ERR_STATE, "%s: this layer cannot carry a fused residual conv", sp.name.c_str());
    const float *xw = nullptr, *xb = nullptr;
    if ((rc = find_w(W, sp.aux_wkey + ".weight", {sp.Cout, cin_real, 1}, &xw))) return rc;
    if ((rc = find_w(W, sp.aux_wkey + ".bias", {sp.Cout}, &xb))) return rc;
    const size_t xplane = (size_t)cout_pad * cin_tot;
    std::vector<__half> xp(2 * xplane, __float2half_rn(0.f));
    int col0 = 0, real0 = 0;
    for (int s = 0; s < sp.nsrc; ++s) {
      for (int ci = 0; ci < sp.src_real_c[s]; ++ci)
        for (int co = 0; co < sp.Cout; ++co) {
          const float v = xw[(size_t)co * cin_real + real0 + ci];
          const __half hi = __float2half_rn(v);
          const size_t o = w_index(0, co, col0 + ci);   // same chunk-major tile layout as the main weights, one tap
          xp[o] = hi;
          xp[xplane + o] = __float2half_rn(v - __half2float(hi));
        }
      col0 += P->acts[sp.src[s]].C;
      real0 += sp.src_real_c[s];
    }
    __half* xdev = nullptr;
    if ((rc = dev_alloc(P, reinterpret_cast<void**>(&xdev), xp.size() * sizeof(__half), false))) return rc;
    DM_CUDA(cudaMemcpy(xdev, xp.data(), xp.size() * sizeof(__half), cudaMemcpyHostToDevice));
    {
      const uint64_t dims[5] = {(uint64_t)j.Cc, (uint64_t)cout_pad, (uint64_t)n_chunks_tot, 1, 2};
      const uint64_t str[4] = {(uint64_t)j.Cc * 2, (uint64_t)cout_pad * j.Cc * 2, (uint64_t)cin_tot * cout_pad * 2,
                               (uint64_t)xplane * 2};
      const uint32_t box[5] = {(uint32_t)j.Cc, (uint32_t)(Nc / 2), 1, 1, 2};
      if ((rc = encode_map(&ly.maps.x_map, xdev, 5, dims, str, box, j.swz))) return rc;
      if (Nc % 32 == 0) {
        const uint32_t qbox[5] = {(uint32_t)j.Cc, (uint32_t)(Nc / 4), 1, 1, 1};
        if ((rc = encode_map(&ly.maps.xq_map, xdev, 5, dims, str, qbox, j.swz))) return rc;
      }
    }
    std::vector<float> hxb(cout_pad, 0.f);
    for (int c = 0; c < sp.Cout; ++c) hxb[c] = xb[c];
    float* dxb;
    if ((rc = upload(P, hxb, &dxb))) return rc;
    int xi;
    if ((rc = new_act(P, cout_pad, Lout, &xi))) return rc;
    {
      const Act& o = P->acts[xi];
      const uint64_t dims[5] = {(uint64_t)o.C, kGroupSamples, (uint64_t)o.L, (uint64_t)P->G, 2};
      const uint64_t str[4] = {(uint64_t)o.C * 2, (uint64_t)o.C * 2 * kGroupSamples,
                               (uint64_t)o.C * 2 * kGroupSamples * o.L, (uint64_t)o.plane * 2};
      const uint32_t box[5] = {(uint32_t)j.st_cw, (uint32_t)j.S, (uint32_t)j.st_lbox, 1, 1};
      if ((rc = encode_map(&ly.maps.p_map, o.p, 5, dims, str, box, j.st_cw == 64 ? 2 : 4))) return rc;
    }
    j.aux = 1;
    j.aux_shift = j.halo_lo;
    j.aux_col0 = j.nphase * (1 + j.acc2) * mt_acc * j.ncols;
    j.aux_bias = dxb;
    ly.aux_act = xi;
    if (aux_idx) *aux_idx = xi;
  }
  // 1: both issuer warps share every tap (k-steps split); 2: whole taps alternate between them (conv_kernel.cu, mma_unit)
  j.mma_mode = j.acc2;
#ifndef DM_MODE2
#define DM_MODE2 2
#endif
#ifndef DM_MODE2_CYCLES
#define DM_MODE2_CYCLES 600
#endif
  if (j.acc2 && j.ntaps >= 2) {
    // pipe time of one tap's MMAs (cycles, MAC floor of the pair MMA) against the issuing warp's per-tap bookkeeping
    const int mma_cycles = j.Mt * j.ksteps * 3 * ((j.mrows * Nc) / 256);
    if (DM_MODE2 == 1 || (DM_MODE2 == 2 && mma_cycles < DM_MODE2_CYCLES)) j.mma_mode = 2;
  }
  // ---- resident link: read the previous layer's staged tile in place (fused loop only) ----
  if (!P->layers.empty()) {
    Layer& pv = P->layers.back();
    const ConvJob& pj = pv.job;
    // (L-split: the tile is one half of the positions; the gap towards the other half receives ITS edge rows from the twin
    // CTA through distributed shared memory, conv_kernel.cu.  A padded half keeps garbage rows behind its valid ones: no link.)
    const int halo_hi_rows = (j.box_l - j.halo_lo - (Lin >> ls)) * j.S;
    const bool ok = !(ls && (pv.ls_padded || ls_padded)) && pv.kind == JOB_CONV && pj.st_cw != 0 && !pj.st_wave && pv.out_act == sp.src[0] && sp.nsrc == 1 && pj.S == 8 && j.S == 8 && pv.slices == 1 &&
                    ly.slices == 1 && pj.st_cw == j.Cc && j.src_chunks[0] == pj.st_nq && j.n_ltiles == 1 &&
                    pj.Lout * pj.S == pj.st_rows && j.halo_lo * j.S * j.row_bytes <= kStageGap &&
                    halo_hi_rows * j.row_bytes <= kStageGap;
    if (ok) {
      const int pv_bytes = 2 * pj.st_nq * pj.st_blk_stride;
      const int w_off = round_up(pv_bytes, 1024);
      int nwr = (budget - w_off) / j.w_stage_bytes;
      if (nwr >= 1) {
        if (nwr < j.n_w_stages) j.n_w_stages = nwr;
        j.a_res = 1;
        j.a_res_off = kStageGap - j.halo_lo * j.S * j.row_bytes;
        j.a_res_chunk = pj.st_blk_stride;
        j.a_res_plane = pj.st_nq * pj.st_blk_stride;
        j.w_off_res = w_off;
        pv.job.next_res = 1;
        const size_t need = (size_t)w_off + (size_t)j.n_w_stages * j.w_stage_bytes + 1024;
        if (ly.smem < need) ly.smem = need;
      }
    }
  }
  if (j.mma_mode == 2) {   // alternating taps need an even ring: a stage then always serves the same issuer warp
    if (j.n_w_stages % 2) j.n_w_stages -= 1;
    if (j.n_w_stages < 2) { j.n_w_stages = 1; j.mma_mode = 1; }
  }
  if (!P->layers.empty() && P->layers.back().kind == JOB_CONV)
    P->layers.back().job.pre_next = prefetchable_w_loads(P->layers.back().job, j);
  P->layers.push_back(ly);
  return DM_OK;
}

}  // namespace dm

// =====================================================================================================
// U-Net launch list (TemporalUnet.__init__/forward, diffuser/models/temporal.py:51-165)
// =====================================================================================================
namespace dm {

struct BlockOut { int act; };

// ResidualTemporalBlock (temporal.py:18-46): conv5+GN+Mish (+time bias) -> conv5+GN+Mish (+ residual)
static int add_resblock(dm_plan* P, const WeightMap& W, const std::string& key, const int* src, const int* src_real,
                        int nsrc, int cin_real, int Cout, int* tb_cursor, int* out_act) {
  int rc, mid = -1, res = -1, out = -1;
  // The 1x1 residual conv shares blocks.0's input: it rides in blocks.0's unit as an auxiliary accumulator set
  // (one extra weight tile per channel chunk, bias-only epilogue).  When TMEM has no room for the second set
  // (very long horizons) it stays a layer of its own, ordered first so that blocks.0 -> blocks.1 remain adjacent.
  ConvSpec c0;
  c0.name = key + ".blocks.0";
  c0.wkey = key + ".blocks.0.block.0";
  c0.gnkey = key + ".blocks.0.block.2";
  c0.kind = CONV_K5;
  c0.nsrc = nsrc;
  for (int s = 0; s < nsrc; ++s) { c0.src[s] = src[s]; c0.src_real_c[s] = src_real[s]; }
  c0.Cout = Cout;
  c0.tb_off = *tb_cursor;
  bool fused_res = false;
  if (cin_real != Cout) {
    c0.aux_wkey = key + ".residual_conv";
    const size_t n_alloc = P->allocs.size(), n_acts = P->acts.size();
    rc = add_conv(P, W, c0, &mid, &res);
    if (rc == DM_OK) {
      fused_res = true;
    } else if (rc == DM_ERR_STATE) {     // does not fit: undo partial allocations of the attempt and fall back
      while (P->allocs.size() > n_alloc) { cudaFree(P->allocs.back()); P->allocs.pop_back(); }
      P->acts.resize(n_acts);
      c0.aux_wkey.clear();
    } else {
      return rc;
    }
  }
  if (!fused_res) {
    if (cin_real != Cout) {
      ConvSpec cr;
      cr.name = key + ".residual_conv";
      cr.wkey = key + ".residual_conv";
      cr.kind = CONV_K1;
      cr.nsrc = nsrc;
      for (int s = 0; s < nsrc; ++s) { cr.src[s] = src[s]; cr.src_real_c[s] = src_real[s]; }
      cr.Cout = Cout;
      if ((rc = add_conv(P, W, cr, &res))) return rc;
    } else {
      res = src[0];  // nn.Identity
    }
    if ((rc = add_conv(P, W, c0, &mid))) return rc;
  }

  ConvSpec c1;
  c1.name = key + ".blocks.1";
  c1.wkey = key + ".blocks.1.block.0";
  c1.gnkey = key + ".blocks.1.block.2";
  c1.kind = CONV_K5;
  c1.src[0] = mid;
  c1.src_real_c[0] = Cout;
  c1.Cout = Cout;
  c1.res = res;
  if ((rc = add_conv(P, W, c1, &out))) return rc;
  *tb_cursor += Cout;
  *out_act = out;
  return DM_OK;
}

// Residual(PreNorm(dim, LinearAttention(dim))) (helpers.py:75-131; temporal.py:85,94,103), present when the state dict
// carries `<key>.fn.fn.to_qkv.weight`.  Weights are uploaded transposed (see attn_kernel.cu).
static bool has_attention_weights(const WeightMap& W, const std::string& key) {
  return W.find(key + ".fn.fn.to_qkv.weight") != W.end();
}

// One CUDA-core stage of the attention block (attn_stages.cuh): a job of the layer list that reads activation `src` and
// writes a new activation of outC channels.
static int add_attn_stage(dm_plan* P, const std::string& name, int kind, int src, int outC, const float* g, const float* b,
                          int L_if_no_src, int* out_act) {
  const int L = src >= 0 ? P->acts[src].L : L_if_no_src;
  int out, rc;
  if ((rc = new_act(P, outC, L, &out))) return rc;
  Layer ly;
  ly.kind = kind;
  ly.hidden = true;
  ly.name = name;
  ly.out_act = out;
  ly.src_acts[0] = src;
  memset(&ly.job, 0, sizeof(ly.job));
  memset(&ly.maps, 0, sizeof(ly.maps));
  ConvJob& j = ly.job;
  j.kind = (unsigned char)kind;
  j.S = 8;
  j.Lj = j.Lout = j.Lfull_out = j.gn_L = L;
  j.Cout = outC;
  j.B = P->B;
  j.slices = 1;
  j.n_rounds = 1;
  j.tb_off = -1;
  if (src >= 0) { const Act& in = P->acts[src]; j.res = in.p; j.res_C = in.C; j.res_plane_stride = in.plane; }
  j.out = P->acts[out].p; j.out_C = outC; j.out_plane_stride = P->acts[out].plane;
  j.gamma = g; j.beta = b;
  ly.smem = kind == JOB_ATTN_CORE ? kCoreSmemBytes + 1024 : 0;
  P->layers.push_back(ly);
  *out_act = out;
  return DM_OK;
}

// Residual(PreNorm(dim, LinearAttention(dim))) (helpers.py:75-131; temporal.py:85,94,103) as four jobs: LayerNorm stage,
// to_qkv 1x1 conv (tensor pipe), softmax/context stage, to_out 1x1 conv with bias and the block's residual (tensor pipe).
// The last one carries the block's name; the first three are hidden from dm_plan_layer_info.
static int add_attention(dm_plan* P, const WeightMap& W, const std::string& key, int src, int C, int* out_act) {
  const Act in = P->acts[src];
  if (in.C != C) return fail(DM_ERR_ARG, "internal: attention input %s is channel padded", key.c_str());
  const float *g, *b;
  int rc;
  if ((rc = find_w(W, key + ".fn.norm.g", {1, C, 1}, &g))) return rc;
  if ((rc = find_w(W, key + ".fn.norm.b", {1, C, 1}, &b))) return rc;
  float *dg, *db;
  if ((rc = upload(P, std::vector<float>(g, g + C), &dg))) return rc;
  if ((rc = upload(P, std::vector<float>(b, b + C), &db))) return rc;
  int xn, qkv, o, y;
  if ((rc = add_attn_stage(P, key + ".norm", JOB_ATTN_NORM, src, C, dg, db, -1, &xn))) return rc;
  ConvSpec cq;
  cq.name = key + ".to_qkv";
  cq.wkey = key + ".fn.fn.to_qkv";
  cq.kind = CONV_K1;
  cq.src[0] = xn;
  cq.src_real_c[0] = C;
  cq.Cout = 3 * kAttnHid;
  cq.no_bias = true;
  cq.hidden = true;
  if ((rc = add_conv(P, W, cq, &qkv))) return rc;
  if ((rc = add_attn_stage(P, key + ".core", JOB_ATTN_CORE, qkv, kAttnHid, nullptr, nullptr, -1, &o))) return rc;
  ConvSpec co;
  co.name = key;
  co.wkey = key + ".fn.fn.to_out";
  co.kind = CONV_K1;
  co.src[0] = o;
  co.src_real_c[0] = kAttnHid;
  co.Cout = C;
  co.res = src;
  if ((rc = add_conv(P, W, co, &y))) return rc;
  P->has_attention = true;
  *out_act = y;
  return DM_OK;
}

// Fused-loop workspace.  Every layer output has an activation of its own (dm_plan_read_layer and the per-layer launches use
// those), 40 x 8.4 MB at the bench shape: far beyond L2, so every tile written is evicted to DRAM before its buffer comes
// round again.  Inside the fused loop a buffer is dead after its last reader of the step, so the loop's copies of the
// descriptors (P->loop_jobs, the device copy of the tensor maps) point activations whose lifetimes do not overlap at the SAME
// memory: the step's working set becomes a handful of buffers that stay in L2.  Lifetimes are by layer index: a layer's
// outputs never share with its own inputs, and a buffer is only handed on after the layer that last reads it (any reader
// from global memory follows a completed store and a barrier, so a later store to the same bytes is ordered behind both).
static int alias_loop_workspace(dm_plan* P, std::vector<ConvMaps>& maps) {
  ReplaceAddressFn replace = get_replace_address();
  if (!replace) return DM_OK;                     // old driver: keep one buffer per activation
  const int n = (int)P->layers.size(), na = (int)P->acts.size();
  std::vector<int> last_use(na, -1);
  for (int i = 0; i < n; ++i) {
    const Layer& ly = P->layers[i];
    for (int s = 0; s < 2; ++s) if (ly.src_acts[s] >= 0) last_use[ly.src_acts[s]] = i;
    if (ly.res_act >= 0) last_use[ly.res_act] = i;
  }
  std::vector<__half*> alias(na);
  for (int a = 0; a < na; ++a) alias[a] = P->acts[a].p;
  std::map<long long, std::vector<__half*>> free_bufs;   // by plane size
  size_t distinct = 0, pooled = 0;
  std::vector<__half*> seen;
  for (int i = 0; i < n; ++i) {
    const Layer& ly = P->layers[i];
    const int defs[2] = {ly.out_act, ly.aux_act};
    for (int d = 0; d < 2; ++d) {
      const int a = defs[d];
      if (a < 0 || a == P->xs_act) continue;
      std::vector<__half*>& fb = free_bufs[P->acts[a].plane];
      if (!fb.empty()) { alias[a] = fb.back(); fb.pop_back(); }
      distinct += (size_t)P->acts[a].plane * 4;
      if (std::find(seen.begin(), seen.end(), alias[a]) == seen.end()) { seen.push_back(alias[a]); pooled += (size_t)P->acts[a].plane * 4; }
    }
    for (int a = 0; a < na; ++a) {
      if (a == P->xs_act) continue;
      const bool def_here = a == ly.out_act || a == ly.aux_act;
      if (last_use[a] == i || (def_here && last_use[a] < i)) free_bufs[P->acts[a].plane].push_back(alias[a]);
    }
  }
  P->loop_workspace_bytes = pooled;
  P->layer_workspace_bytes = distinct;
  auto swap = [&](CUtensorMap* m, int a) -> int {
    if (a < 0 || alias[a] == P->acts[a].p) return DM_OK;
    const CUresult r = replace(m, alias[a]);
    return r == CUDA_SUCCESS ? DM_OK : fail(DM_ERR_CUDA, "cuTensorMapReplaceAddress failed (%d)", (int)r);
  };
  for (int i = 0; i < n; ++i) {
    const Layer& ly = P->layers[i];
    ConvJob& j = P->loop_jobs[i];
    int rc;
    if (ly.kind == JOB_CONV) {
      for (int s = 0; s < j.nsrc; ++s) if ((rc = swap(&maps[i].a_map[s], ly.src_acts[s]))) return rc;
      if (j.st_cw != 0 && ly.out_act >= 0 && (rc = swap(&maps[i].o_map, ly.out_act))) return rc;
      if (j.st_cw != 0 && ly.res_act >= 0 && (rc = swap(&maps[i].r_map, ly.res_act))) return rc;
      if (j.aux && (rc = swap(&maps[i].p_map, ly.aux_act))) return rc;
      if (ly.res_act >= 0) j.res = alias[ly.res_act];
    } else if (ly.src_acts[0] >= 0) {
      j.res = alias[ly.src_acts[0]];
    }
    if (ly.out_act >= 0) j.out = alias[ly.out_act];
  }
  return DM_OK;
}

static int build_unet(dm_plan* P, const WeightMap& W) {
  const int nlev = (int)P->mults.size();
  std::vector<int> dims;
  dims.push_back(P->T);
  for (int m : P->mults) dims.push_back(P->dim * m);
  int rc;

  // time-bias row layout: blocks in module order downs, mid, ups
  int tb_total = 0;
  for (int i = 0; i < nlev; ++i) tb_total += 2 * dims[i + 1];
  tb_total += 2 * dims[nlev];
  for (int i = nlev - 1; i >= 1; --i) tb_total += 2 * dims[i];
  P->TB = tb_total;

  // input operand planes: transition_dim padded to a multiple of 64 channels
  const int xs_C = round_up(P->T, 64);
  if ((rc = new_act(P, xs_C, P->H, &P->xs_act))) return rc;
  if ((rc = dev_alloc(P, reinterpret_cast<void**>(&P->model_out), (size_t)P->B * P->H * P->T * sizeof(float)))) return rc;

  int tb = 0;
  int x = P->xs_act, x_real = P->T;
  std::vector<int> skips, skip_c;
  std::vector<float> wcat, bcat;
  auto add_tb = [&](const std::string& key, int Cout) -> int {
    const float *w = nullptr, *b = nullptr;
    int r;
    if ((r = find_w(W, key + ".time_mlp.1.weight", {Cout, P->dim}, &w))) return r;
    if ((r = find_w(W, key + ".time_mlp.1.bias", {Cout}, &b))) return r;
    wcat.insert(wcat.end(), w, w + (size_t)Cout * P->dim);
    bcat.insert(bcat.end(), b, b + Cout);
    return DM_OK;
  };

  for (int i = 0; i < nlev; ++i) {
    const int cout = dims[i + 1];
    const std::string k = "downs." + std::to_string(i);
    int o;
    if ((rc = add_tb(k + ".0", cout))) return rc;
    if ((rc = add_resblock(P, W, k + ".0", &x, &x_real, 1, x_real, cout, &tb, &o))) return rc;
    x = o; x_real = cout;
    if ((rc = add_tb(k + ".1", cout))) return rc;
    if ((rc = add_resblock(P, W, k + ".1", &x, &x_real, 1, x_real, cout, &tb, &o))) return rc;
    x = o;
    if (has_attention_weights(W, k + ".2")) {
      if ((rc = add_attention(P, W, k + ".2", x, cout, &o))) return rc;
      x = o;
    }
    skips.push_back(x);
    skip_c.push_back(cout);
    if (i < nlev - 1) {
      ConvSpec d;
      d.name = k + ".3.conv";
      d.wkey = k + ".3.conv";
      d.kind = CONV_DOWN;
      d.src[0] = x;
      d.src_real_c[0] = cout;
      d.Cout = cout;
      if ((rc = add_conv(P, W, d, &o))) return rc;
      x = o;
    }
  }
  {
    const int c = dims[nlev];
    int o;
    if ((rc = add_tb("mid_block1", c))) return rc;
    if ((rc = add_resblock(P, W, "mid_block1", &x, &x_real, 1, c, c, &tb, &o))) return rc;
    x = o;
    if (has_attention_weights(W, "mid_attn")) {
      if ((rc = add_attention(P, W, "mid_attn", x, c, &o))) return rc;
      x = o;
    }
    if ((rc = add_tb("mid_block2", c))) return rc;
    if ((rc = add_resblock(P, W, "mid_block2", &x, &x_real, 1, c, c, &tb, &o))) return rc;
    x = o;
  }
  // ups: for (dim_in, dim_out) in reversed(in_out[1:]) (temporal.py:96-109)
  for (int u = 0; u < nlev - 1; ++u) {
    const int lev = nlev - 1 - u;          // in_out index
    const int dim_in = dims[lev], dim_out = dims[lev + 1];
    const std::string k = "ups." + std::to_string(u);
    const int skip = skips.back();
    skips.pop_back();
    skip_c.pop_back();
    int srcs[2] = {x, skip};
    int reals[2] = {dim_out, dim_out};
    int o;
    if ((rc = add_tb(k + ".0", dim_in))) return rc;
    if ((rc = add_resblock(P, W, k + ".0", srcs, reals, 2, 2 * dim_out, dim_in, &tb, &o))) return rc;
    x = o; x_real = dim_in;
    if ((rc = add_tb(k + ".1", dim_in))) return rc;
    if ((rc = add_resblock(P, W, k + ".1", &x, &x_real, 1, dim_in, dim_in, &tb, &o))) return rc;
    x = o;
    if (has_attention_weights(W, k + ".2")) {
      if ((rc = add_attention(P, W, k + ".2", x, dim_in, &o))) return rc;
      x = o;
    }
    ConvSpec t;
    t.name = k + ".3.conv";
    t.wkey = k + ".3.conv";
    t.kind = CONV_T4;
    t.src[0] = x;
    t.src_real_c[0] = dim_in;
    t.Cout = dim_in;
    if ((rc = add_conv(P, W, t, &o))) return rc;
    x = o;
  }
  {
    ConvSpec f0;
    f0.name = "final_conv.0";
    f0.wkey = "final_conv.0.block.0";
    f0.gnkey = "final_conv.0.block.2";
    f0.kind = CONV_K5;
    f0.src[0] = x;
    f0.src_real_c[0] = P->dim;
    f0.Cout = P->dim;
    int o;
    if ((rc = add_conv(P, W, f0, &o))) return rc;
    ConvSpec f1;
    f1.name = "final_conv.1";
    f1.wkey = "final_conv.1";
    f1.kind = CONV_K1;
    f1.src[0] = o;
    f1.src_real_c[0] = P->dim;
    f1.Cout = P->T;
    f1.want_act_out = false;
    f1.f32_out = true;
    int o2;
    if ((rc = add_conv(P, W, f1, &o2))) return rc;
  }
  P->layers.back().job.pre_next = prefetchable_w_loads(P->layers.back().job, P->layers.front().job);
  for (size_t i = 0; i + 1 < P->layers.size(); ++i)
      P->layers[i].job.sync_local = !P->lsplit && P->layers[i].job.S == 8 && P->layers[i + 1].job.S == 8;
  if (tb != P->TB) return fail(DM_ERR_ARG, "internal: time-bias layout mismatch (%d vs %d)", tb, P->TB);
  // Scratch for the GroupNorm layers that run in TMEM rounds: [G][Lj][16][cout_pad] fp32, shared by all of them.
  {
    size_t need = 0;
    for (Layer& ly : P->layers)
      if (ly.kind == 0 && ly.job.n_rounds > 1 && ly.job.epi == EPI_GN_MISH) {
        const size_t n = (size_t)P->G * ly.job.Lj * kGroupSamples * round_up(ly.job.Cout, 32);
        if (n > need) need = n;
      }
    if (need) {
      float* raw = nullptr;
      if ((rc = dev_alloc(P, reinterpret_cast<void**>(&raw), need * sizeof(float), false))) return rc;
      for (Layer& ly : P->layers)
        if (ly.kind == 0 && ly.job.n_rounds > 1 && ly.job.epi == EPI_GN_MISH) {
          ly.job.raw = raw;
          ly.job.raw_C = round_up(ly.job.Cout, 32);
        }
    }
  }
  // Halo gaps of a staged tile stay zero when the bytes were last written by the previous layer's tile of the same geometry
  // (this layer read that tile in place, its weights ring lies beyond it, residual / auxiliary tiles only touch data rows).
  for (size_t i = 1; i < P->layers.size(); ++i) {
    ConvJob& j = P->layers[i].job;
    const ConvJob& pj = P->layers[i - 1].job;
    if (P->layers[i].kind != 0 || P->layers[i - 1].kind != 0) continue;
#ifndef DM_NO_GAPS_CLEAN
    j.gaps_clean = !P->lsplit && j.next_res && j.a_res && pj.next_res && pj.st_cw == j.st_cw && pj.st_nq == j.st_nq &&
                   pj.st_blk_stride == j.st_blk_stride && pj.st_rows == j.st_rows && j.f32_stride == 0;
#endif
  }
  // An output tile whose only reader is the next layer's in-place (resident) read never has to reach global memory.
  for (size_t i = 0; i + 1 < P->layers.size(); ++i) {
    Layer& ly = P->layers[i];
    if (ly.kind != 0 || !ly.job.next_res || ly.out_act < 0) continue;
    int readers = 0;
    for (size_t k = 0; k < P->layers.size(); ++k) {
      const Layer& c = P->layers[k];
      for (int s = 0; s < 2; ++s) readers += c.src_acts[s] == ly.out_act;
      readers += c.res_act == ly.out_act;
    }
    if (readers == 1 && P->layers[i + 1].src_acts[0] == ly.out_act) ly.job.no_store = 1;
  }

  // time MLP parameters
  const float *w1, *b1, *w2, *b2;
  if ((rc = find_w(W, "time_mlp.1.weight", {4 * P->dim, P->dim}, &w1))) return rc;
  if ((rc = find_w(W, "time_mlp.1.bias", {4 * P->dim}, &b1))) return rc;
  if ((rc = find_w(W, "time_mlp.3.weight", {P->dim, 4 * P->dim}, &w2))) return rc;
  if ((rc = find_w(W, "time_mlp.3.bias", {P->dim}, &b2))) return rc;
  float *dw1, *db1, *dw2, *db2, *dwc, *dbc;
  const int D = P->dim;
  if ((rc = upload(P, std::vector<float>(w1, w1 + 4 * D * D), &dw1))) return rc;
  if ((rc = upload(P, std::vector<float>(b1, b1 + 4 * D), &db1))) return rc;
  if ((rc = upload(P, std::vector<float>(w2, w2 + 4 * D * D), &dw2))) return rc;
  if ((rc = upload(P, std::vector<float>(b2, b2 + D), &db2))) return rc;
  if ((rc = upload(P, wcat, &dwc))) return rc;
  if ((rc = upload(P, bcat, &dbc))) return rc;
  P->tmlp = TimeMlpParams{D, P->TB, dw1, db1, dw2, db2, dwc, dbc};
  if ((rc = dev_alloc(P, reinterpret_cast<void**>(&P->seed_dev), sizeof(unsigned long long)))) return rc;
  {
    const int n = (int)P->layers.size();
    if (n > kMaxLoopJobs) return fail(DM_ERR_ARG, "U-Net has %d layer launches; the fused loop supports %d", n, kMaxLoopJobs);
    std::vector<ConvMaps> maps(n);
    P->loop_smem = 0;
    P->loop_jobs.resize(n);
    for (int i = 0; i < n; ++i) {
      maps[i] = P->layers[i].maps;
      P->loop_jobs[i] = P->layers[i].job;
      if (P->layers[i].smem > P->loop_smem) P->loop_smem = P->layers[i].smem;
    }
    if ((rc = alias_loop_workspace(P, maps))) return rc;
    // Weight multicast (conv_kernel.cu, producer_unit): clusters of four CTAs = two groups whose pairs share every weight
    // tile.  Common kernel shape only, an even number of groups, and at least one full cluster's worth of them.
#ifndef DM_MCAST_MAX_H
#define DM_MCAST_MAX_H 128
#endif
#ifndef DM_NO_MCAST
    // Measured (A/B, ms per pass): B = 1024, H = 64: 43.2 -> 40.0; B = 2048, H = 128: 158.2 -> 153.0.  The multicast form costs
    // four small TMA operations per tile instead of two, and at long horizons the first levels re-stream their 2 KB tiles per
    // l-tile: B = 256, H = 480 (L-split) 172.9 -> 187.7, and the L-split layouts at H = 128 / 256 gain under 1 %.
    P->mcast = !P->has_attention && !P->lsplit && P->G % 2 == 0 && P->H <= DM_MCAST_MAX_H;
    for (int i = 0; i < n; ++i) {
      const Layer& ly = P->layers[i];
      if (ly.kind != JOB_CONV || !ly.mc_ok) P->mcast = false;
    }
#endif
    if (P->mcast || P->lsplit) {
      // Both layouts launch clusters of four; how many of them the device holds at once is decided by GPC boundaries and
      // disabled SMs (33 on a 148-SM B200, against 74 pairs).  The split must fit in ONE wave (it exists to fill idle SMs);
      // multicast must not need more waves than the independent pairs would.
      int fit4 = 0, fit2 = 0;
      if (max_resident_clusters(4, P->loop_smem, &fit4) != cudaSuccess || fit4 < 1) { cudaGetLastError(); fit4 = 0; }
      if (max_resident_clusters(2, P->loop_smem, &fit2) != cudaSuccess || fit2 < 1) { cudaGetLastError(); fit2 = 1; }
      if (P->lsplit && fit4 < P->G)
        return fail(DM_ERR_STATE, "the device holds %d clusters of four, the split plan needs %d", fit4, P->G);
      if (P->mcast && (fit4 == 0 || (P->G / 2 + fit4 - 1) / fit4 > (P->G + fit2 - 1) / fit2)) P->mcast = false;
    }
    if (P->mcast)
      for (int i = 0; i < n; ++i) {
        P->loop_jobs[i].mc = 1;           // (every job or none: the kernel variant is chosen by the first job's flag)
        // The other pair's producer fills THIS CTA's weight stages for the next layer as soon as ITS roles are through the
        // barrier: a block barrier is enough only where those stages lie beyond this layer's staging (what pre_next says).
        if (P->loop_jobs[i].pre_next == 0) P->loop_jobs[i].sync_local = 0;
      }
    if ((rc = dev_alloc(P, reinterpret_cast<void**>(&P->maps_dev), n * sizeof(ConvMaps), false))) return rc;
    DM_CUDA(cudaMemcpy(P->maps_dev, maps.data(), n * sizeof(ConvMaps), cudaMemcpyHostToDevice));
  }
  if ((rc = dev_alloc(P, reinterpret_cast<void**>(&P->tb_rows), (size_t)P->G * kGroupSamples * P->TB * sizeof(float))))
    return rc;
  return DM_OK;
}

// Launch every layer.  tb: time-bias rows; tb_stride 0 = one row for all samples (uniform t).
static int run_unet(dm_plan* P, const float* tb, int tb_stride, float* out_f32, cudaStream_t s) {
  for (Layer& ly : P->layers) {
    if (ly.kind != JOB_CONV) {
      DM_CUDA(launch_attention_stage(ly.job, P->G, s));
      P->last_launches++;
      continue;
    }
    ConvJob j = ly.job;
    if (ly.tb_off >= 0) {
      j.tbias = tb + ly.tb_off;
      j.tbias_stride = tb_stride;
    }
    if (j.out_f32 == P->model_out) j.out_f32 = out_f32;
    j.a_res = 0;       // per-layer launches always go through global memory
    j.next_res = 0;
    j.no_store = 0;
    j.gaps_clean = 0;
    j.pre_next = 0;
    j.sync_local = 0;
    DM_CUDA(launch_conv_job(ly.maps, j, P->G, ly.slices, ly.smem, s));
    P->last_launches++;
#ifdef DM_DEV_DEBUG
    {   // debugging aid (compile with -DDM_DEV_DEBUG): find the layer whose kernel faults
      const cudaError_t e = cudaStreamSynchronize(s);
      fprintf(stderr, "[dm debug] %-28s mode %d S %d Mt %d ltiles %d Nc %d slices %d chunks %d+%d taps %d aux %d : %s\n", ly.name.c_str(),
              j.mma_mode, j.S, j.Mt, j.n_ltiles, j.Nc, j.slices, j.src_chunks[0], j.nsrc > 1 ? j.src_chunks[1] : 0, j.ntaps, j.aux,
              cudaGetErrorString(e));
      if (e != cudaSuccess) return fail(DM_ERR_CUDA, "layer %s failed", ly.name.c_str());
    }
#endif
  }
  return DM_OK;
}

}  // namespace dm

// =====================================================================================================
// C ABI
// =====================================================================================================
using namespace dm;

extern "C" const char* dm_last_error(void) { return last_error(); }

extern "C" int dm_device_error_code(unsigned int* code) {
  if (!code) return fail(DM_ERR_ARG, "null argument");
  DM_CUDA(read_device_error(code));
  return DM_OK;
}

extern "C" int dm_plan_create(const dm_tensor_t* weights, int n_weights, int B, int H, int T, int dim,
                              const int* dim_mults, int n_mults, dm_plan_t** plan) {
  if (!weights || !plan || !dim_mults) return fail(DM_ERR_ARG, "null argument");
  if (B < 1 || T < 1 || dim < 32 || dim % 32) return fail(DM_ERR_ARG, "need B >= 1, T >= 1, dim a multiple of 32");
  if (n_mults < 2 || n_mults > 6) return fail(DM_ERR_ARG, "need 2..6 resolution levels");
  const int div = 1 << (n_mults - 1);
  if (H < div || H % div)
    return fail(DM_ERR_ARG, "horizon %d must be a positive multiple of %d (temporal.py:156 concat would fail)", H, div);
  int rc = check_device();
  if (rc) return rc;
  WeightMap W;
  for (int i = 0; i < n_weights; ++i) {
    if (!weights[i].name || !weights[i].data || weights[i].ndim < 1 || weights[i].ndim > 4)
      return fail(DM_ERR_WEIGHT, "weight entry %d is malformed", i);
    HostTensor t;
    t.data = weights[i].data;
    t.shape.assign(weights[i].shape, weights[i].shape + weights[i].ndim);
    W[weights[i].name] = t;
  }
  // L-split (conv_job.h): when the 16-sample groups would leave more than half of the SMs idle and the horizon is long
  // enough for every level to split into whole MMA tiles, two CTA pairs share each group.  Layers that cannot be split
  // (DM_ERR_STATE from add_conv) or an attention model send the build back to the unsplit layout.
  const int G = (B + kGroupSamples - 1) / kGroupSamples;
  int n_sm = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
#ifndef DM_LSPLIT_MIN_H
#define DM_LSPLIT_MIN_H 128
#endif
#ifdef DM_NO_LSPLIT
  bool try_split = false;
#else
  bool try_split = 4 * G <= n_sm && H >= DM_LSPLIT_MIN_H && (H >> (n_mults - 1)) % 2 == 0;
#endif
  dm_plan* P = nullptr;
  for (;;) {
    P = new dm_plan();
    P->B = B; P->H = H; P->T = T; P->dim = dim;
    P->G = G;
    P->lsplit = try_split;
    P->mults.assign(dim_mults, dim_mults + n_mults);
    P->device = dev;
    if (cudaEventCreateWithFlags(&P->steps_event, cudaEventDisableTiming) != cudaSuccess) { delete P; return fail(DM_ERR_CUDA, "event create failed"); }
    rc = build_unet(P, W);
    if (try_split && (rc == DM_ERR_STATE || (rc == DM_OK && P->has_attention))) {
      dm_plan_free(P);
      try_split = false;
      continue;
    }
    break;
  }
  if (rc) { dm_plan_free(P); return rc; }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { dm_plan_free(P); return fail(DM_ERR_CUDA, "plan upload failed: %s", cudaGetErrorString(e)); }
  *plan = P;
  return DM_OK;
}

extern "C" void dm_plan_free(dm_plan_t* P) {
  if (!P) return;
  if (P->steps_host) cudaFreeHost(P->steps_host);
  if (P->steps_event) cudaEventDestroy(P->steps_event);
  for (void* p : P->allocs) cudaFree(p);
  delete P;
}

static unsigned long long fnv1a(unsigned long long h, const void* data, size_t bytes) {
  const unsigned char* p = static_cast<const unsigned char*>(data);
  for (size_t i = 0; i < bytes; ++i) { h ^= p[i]; h *= 1099511628211ull; }
  return h;
}

static void free_tracked(dm_plan* P, void* p) {
  if (!p) return;
  for (size_t i = 0; i < P->allocs.size(); ++i)
    if (P->allocs[i] == p) { P->allocs.erase(P->allocs.begin() + i); break; }
  cudaFree(p);
}

extern "C" int dm_plan_set_schedule(dm_plan_t* P, const dm_schedule_t* s) {
  if (!P || !s || s->n_timesteps < 1) return fail(DM_ERR_ARG, "bad schedule");
  int rc = check_plan_device(P);
  if (rc) return rc;
  const int n = s->n_timesteps;
  std::vector<float> tv(n);
  for (int i = 0; i < n; ++i) tv[i] = s->time_values ? s->time_values[i] : (float)i;
  unsigned long long h = fnv1a(14695981039346656037ull, &n, sizeof(n));
  const float* arrs[5] = {s->sqrt_recip_alphas_cumprod, s->sqrt_recipm1_alphas_cumprod, s->posterior_mean_coef1,
                          s->posterior_mean_coef2, s->posterior_log_variance_clipped};
  for (const float* a : arrs) h = fnv1a(h, a, (size_t)n * sizeof(float));
  h = fnv1a(h, tv.data(), (size_t)n * sizeof(float));
  // step table staging sized for the longest schedule seen (one denoising step per entry + the init entry)
  if (n + 1 > P->steps_capacity) {
    if (P->steps_pending) { DM_CUDA(cudaEventSynchronize(P->steps_event)); P->steps_pending = false; }
    DM_CUDA(cudaDeviceSynchronize());   // a running loop kernel may still read the old table
    free_tracked(P, P->steps_dev);
    P->steps_dev = nullptr;
    if (P->steps_host) { cudaFreeHost(P->steps_host); P->steps_host = nullptr; }
    if ((rc = dev_alloc(P, reinterpret_cast<void**>(&P->steps_dev), (size_t)(n + 1) * sizeof(StepParams)))) return rc;
    DM_CUDA(cudaMallocHost(reinterpret_cast<void**>(&P->steps_host), (size_t)(n + 1) * sizeof(StepParams)));
    P->steps_capacity = n + 1;
  }
  ScheduleTables* hit = nullptr;
  for (size_t i = 0; i < P->schedules.size(); ++i)
    if (P->schedules[i].hash == h && P->schedules[i].n == n) {
      std::rotate(P->schedules.begin(), P->schedules.begin() + i, P->schedules.begin() + i + 1);   // most recent first
      hit = &P->schedules.front();
      break;
    }
  if (!hit) {
    ScheduleTables e;
    if (P->schedules.size() >= 4) {   // recycle the least recently used entry (nothing on the device may still read it)
      DM_CUDA(cudaDeviceSynchronize());
      free_tracked(P, P->schedules.back().t_table);
      free_tracked(P, P->schedules.back().tb_table);
      P->schedules.pop_back();
    }
    e.hash = h;
    e.n = n;
    e.sr.assign(arrs[0], arrs[0] + n);
    e.srm1.assign(arrs[1], arrs[1] + n);
    e.c1.assign(arrs[2], arrs[2] + n);
    e.c2.assign(arrs[3], arrs[3] + n);
    e.logvar.assign(arrs[4], arrs[4] + n);
    if ((rc = upload(P, tv, &e.t_table))) return rc;
    if ((rc = dev_alloc(P, reinterpret_cast<void**>(&e.tb_table), (size_t)n * P->TB * sizeof(float)))) return rc;
    DM_CUDA(launch_time_bias(P->tmlp, e.t_table, n, e.tb_table, 0));   // legacy stream: ordered before later work
    DM_CUDA(cudaStreamSynchronize(0));
    P->schedules.insert(P->schedules.begin(), std::move(e));
    hit = &P->schedules.front();
  }
  P->sr = hit->sr; P->srm1 = hit->srm1; P->c1 = hit->c1; P->c2 = hit->c2; P->logvar = hit->logvar;
  P->t_table = hit->t_table;
  P->tb_table = hit->tb_table;
  P->n_timesteps = n;
  return DM_OK;
}

extern "C" int dm_unet_forward(dm_plan_t* P, const float* x_dev, const float* t_dev, float* out_dev, dm_stream_t st) {
  if (!P || !x_dev || !t_dev || !out_dev) return fail(DM_ERR_ARG, "null argument");
  if (int drc = check_plan_device(P)) return drc;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  P->last_launches = 0;
  const Act& xs = P->acts[P->xs_act];
  DM_CUDA(launch_pack_input(x_dev, xs.p, xs.plane, xs.C, P->B, P->H, P->T, s));
  DM_CUDA(launch_time_bias(P->tmlp, t_dev, P->B, P->tb_rows, s));
  P->last_launches += 2;
  return run_unet(P, P->tb_rows, P->TB, out_dev, s);
}

static int fill_step(dm_plan* P, StepParams* sp, float* x, const float* model_out, const float* noise,
                     unsigned long long seed, int t, int pe, int clip, const dm_cond_t* cond) {
  memset(sp, 0, sizeof(*sp));
  const Act& xs = P->acts[P->xs_act];
  sp->x = x;
  sp->model_out = model_out;
  sp->noise = noise;
  sp->seed = seed;
  sp->predict_epsilon = pe;
  sp->clip_denoised = clip;
  if (t >= 0) {
    sp->mode = 0;
    sp->t_index = t;
    sp->add_noise = t > 0;   // noise[t == 0] = 0 (diffusion.py:25-26)
    sp->sr = P->sr[t]; sp->srm1 = P->srm1[t]; sp->c1 = P->c1[t]; sp->c2 = P->c2[t];
    sp->sigma = expf(0.5f * P->logvar[t]);
  } else {
    sp->mode = 1;
  }
  if (cond && cond->mask_dev) {
    if (!cond->value_dev) return fail(DM_ERR_ARG, "conditioning mask without values");
    sp->mask = cond->mask_dev;
    sp->value = cond->value_dev;
    sp->mask_bstride = cond->mask_batched ? (long long)P->H * P->T : 0;
    sp->value_bstride = cond->value_batched ? (long long)P->H * P->T : 0;
    if (cond->flags_dev) {
      if (!cond->mask_batched || !cond->value_batched)
        return fail(DM_ERR_ARG, "dm_cond_t::flags_dev needs the full [B,H*T] mask and value tensors");
      sp->cond_flags = cond->flags_dev;
    }
  }
  sp->xs = xs.p;
  sp->xs_plane_stride = xs.plane;
  sp->xs_C = xs.C;
  sp->B = P->B; sp->H = P->H; sp->T = P->T;
  sp->n_elems = (long long)P->B * P->H * P->T;
  if (sp->n_elems % 4) return fail(DM_ERR_ARG, "B*H*T must be a multiple of 4");
  // Philox counter = index of the float4 inside the GLOBAL [B_total, H, T] tensor: a shard that starts at sample
  // `sample_offset` draws exactly the noise a single-GPU run would draw for those rows
  sp->rng_offset = (unsigned long long)P->sample_offset * (unsigned long long)(((long long)P->H * P->T) >> 2);
  return DM_OK;
}

extern "C" int dm_ddpm_step(dm_plan_t* P, float* x_dev, const float* model_out_dev, const float* noise_dev,
                            uint64_t seed, int t, int predict_epsilon, int clip_denoised, const dm_cond_t* cond,
                            dm_stream_t st) {
  if (!P || !x_dev || !model_out_dev) return fail(DM_ERR_ARG, "null argument");
  if (P->n_timesteps == 0) return fail(DM_ERR_STATE, "dm_plan_set_schedule has not been called");
  if (t < 0 || t >= P->n_timesteps) return fail(DM_ERR_ARG, "timestep %d out of range", t);
  if (int drc = check_plan_device(P)) return drc;
  StepParams sp;
  int rc = fill_step(P, &sp, x_dev, model_out_dev, noise_dev, seed, t, predict_epsilon, clip_denoised, cond);
  if (rc) return rc;
  sp.rng_stream = (unsigned)t + 1;
  sp.rng_offset += P->rng_calls++ << 40;
  DM_CUDA(launch_ddpm_step(sp, static_cast<cudaStream_t>(st)));
  P->last_launches = 1;
  return DM_OK;
}

static int enqueue_loop(dm_plan* P, float* x, int have_start, const float* noise_stack, unsigned long long seed,
                        int timesteps, int pe, int clip, const dm_cond_t* cond, float* chain, cudaStream_t s) {
  const long long n = (long long)P->B * P->H * P->T;
  StepParams sp;
  int rc = fill_step(P, &sp, x, nullptr, noise_stack, seed, -1, pe, clip, cond);
  if (rc) return rc;
  sp.have_start = have_start;
  sp.rng_stream = 0;
  sp.seed_ptr = P->seed_dev;
  sp.chain = chain; sp.chain_len = timesteps + 1; sp.chain_slot = 0;
  DM_CUDA(launch_ddpm_step(sp, s));
  P->last_launches++;
  int k = 1;
  for (int i = timesteps - 1; i >= 0; --i, ++k) {
    if ((rc = run_unet(P, P->tb_table + (size_t)i * P->TB, 0, P->model_out, s))) return rc;
    if ((rc = fill_step(P, &sp, x, P->model_out, noise_stack ? noise_stack + (size_t)k * n : nullptr, seed, i, pe, clip,
                        cond)))
      return rc;
    sp.rng_stream = (unsigned)k;
    sp.seed_ptr = P->seed_dev;
    sp.chain = chain; sp.chain_len = timesteps + 1; sp.chain_slot = k;
    DM_CUDA(launch_ddpm_step(sp, s));
    P->last_launches++;
  }
  return DM_OK;
}

extern "C" int dm_sample_loop(dm_plan_t* P, float* x_dev, int have_start, const float* noise_stack_dev, uint64_t seed,
                              int timesteps, int predict_epsilon, int clip_denoised, const dm_cond_t* cond,
                              float* chain_dev, int fused, dm_stream_t st) {
  if (!P || !x_dev) return fail(DM_ERR_ARG, "null argument");
  if (P->n_timesteps == 0) return fail(DM_ERR_STATE, "dm_plan_set_schedule has not been called");
  if (timesteps < 0 || timesteps > P->n_timesteps) return fail(DM_ERR_ARG, "timesteps %d out of range", timesteps);
  if (int drc = check_plan_device(P)) return drc;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  P->last_launches = 0;
  if (P->steps_pending) { DM_CUDA(cudaEventSynchronize(P->steps_event)); P->steps_pending = false; }
  {
    unsigned long long sd = seed;
    DM_CUDA(cudaMemcpyAsync(P->seed_dev, &sd, sizeof(sd), cudaMemcpyHostToDevice, s));
  }
  if (!fused) return enqueue_loop(P, x_dev, have_start, noise_stack_dev, seed, timesteps, predict_epsilon, clip_denoised,
                                  cond, chain_dev, s);
  // ---- persistent cluster kernel: the whole loop is ONE launch ----
  const long long n = (long long)P->B * P->H * P->T;
  StepParams* sp = P->steps_host;
  int rc = fill_step(P, &sp[0], x_dev, nullptr, noise_stack_dev, seed, -1, predict_epsilon, clip_denoised, cond);
  if (rc) return rc;
  sp[0].have_start = have_start;
  sp[0].rng_stream = 0;
  sp[0].seed_ptr = P->seed_dev;
  sp[0].chain = chain_dev; sp[0].chain_len = timesteps + 1; sp[0].chain_slot = 0;
  int k = 1;
  for (int i = timesteps - 1; i >= 0; --i, ++k) {
    if ((rc = fill_step(P, &sp[k], x_dev, P->model_out, noise_stack_dev ? noise_stack_dev + (size_t)k * n : nullptr, seed,
                        i, predict_epsilon, clip_denoised, cond)))
      return rc;
    sp[k].rng_stream = (unsigned)k;
    sp[k].seed_ptr = P->seed_dev;
    sp[k].chain = chain_dev; sp[k].chain_len = timesteps + 1; sp[k].chain_slot = k;
  }
  DM_CUDA(cudaMemcpyAsync(P->steps_dev, sp, (size_t)(timesteps + 1) * sizeof(StepParams), cudaMemcpyHostToDevice, s));
  // the pinned step table is rewritten by the next call: that call waits for THIS copy only (not for the loop kernel
  // behind it), so back-to-back passes queue without a host stall
  DM_CUDA(cudaEventRecord(P->steps_event, s));
  P->steps_pending = true;
  static thread_local LoopParams lp;   // ~19 KB: keep it off the stack
  lp.n_jobs = (int)P->layers.size();
  for (int i = 0; i < lp.n_jobs; ++i) lp.jobs[i] = P->loop_jobs[i];   // (activation pointers of the shared loop workspace)
#ifdef DM_DEV_DEBUG
  {   // debugging aid (compile with -DDM_DEV_DEBUG): the fused loop's job list
    for (int i = 0; i < lp.n_jobs; ++i) {
      const ConvJob& j = lp.jobs[i];
      fprintf(stderr, "[dm plan] %2d %-24s S %2d Mt %d lt %d Nc %3d sl %d ch %d+%d taps %d aux %d mode %d acc2 %d | A st %d x %6d  W st %d x %6d "
              "a_res %d w_off %6d pre_next %d next_res %d local %d cols %d\n", i, P->layers[i].name.c_str(), j.S, j.Mt, j.n_ltiles, j.Nc,
              j.slices, j.src_chunks[0], j.nsrc > 1 ? j.src_chunks[1] : 0, j.ntaps, j.aux, j.mma_mode, j.acc2, j.n_a_stages,
              j.a_stage_bytes, j.n_w_stages, j.w_stage_bytes, j.a_res, j.a_res ? j.w_off_res : j.n_a_stages * j.a_stage_bytes, j.pre_next,
              j.next_res, j.sync_local, j.tmem_cols);
    }
  }
#endif
  lp.maps = P->maps_dev;
  lp.steps = P->steps_dev; lp.n_steps = timesteps; lp.tb_table = P->tb_table; lp.TB = P->TB;
  lp.G = P->G; lp.B = P->B; lp.H = P->H; lp.T = P->T;
  lp.k4_local = !P->lsplit && (!P->mcast || P->layers.back().job.pre_next > 0) && P->layers.front().job.S == 8 &&
                P->layers.back().job.S == 8;
  lp.trace = P->trace_dev; lp.trace_cluster = P->trace_cluster; lp.trace_step = P->trace_step;
  DM_CUDA(launch_sample_loop(lp, P->loop_smem, s));
  P->last_launches = 1;
  return DM_OK;
}

extern "C" int dm_plan_num_layers(const dm_plan_t* P) { return P ? (int)P->layers.size() : 0; }

extern "C" int dm_plan_ctas_per_group(const dm_plan_t* P) { return P ? (P->lsplit ? 4 : 2) : 0; }

extern "C" int dm_plan_layer_info(const dm_plan_t* P, int i, char* name, int name_cap, int* C, int* L) {
  if (!P || i < 0 || i >= (int)P->layers.size()) return fail(DM_ERR_ARG, "layer index out of range");
  const Layer& ly = P->layers[i];
  if (name && name_cap > 0) snprintf(name, name_cap, "%s", ly.name.c_str());
  if (C) *C = ly.out_act >= 0 && !ly.hidden ? ly.job.Cout : 0;
  if (L) *L = ly.kind == 0 ? ly.job.Lfull_out : ly.job.Lout;
  return DM_OK;
}

extern "C" int dm_plan_read_layer(dm_plan_t* P, int i, float* out_dev, dm_stream_t st) {
  if (!P || i < 0 || i >= (int)P->layers.size() || !out_dev) return fail(DM_ERR_ARG, "layer index out of range");
  const Layer& ly = P->layers[i];
  if (ly.out_act < 0) return fail(DM_ERR_ARG, "layer %d has no activation output", i);
  const Act& a = P->acts[ly.out_act];
  // unpack only the real channels: out is [B, Cout, L]
  if (a.C != ly.job.Cout) return fail(DM_ERR_ARG, "layer %d output is channel padded", i);
  DM_CUDA(launch_unpack_act(a.p, a.plane, a.C, a.L, P->B, out_dev, static_cast<cudaStream_t>(st)));
  return DM_OK;
}

extern "C" int dm_plan_set_trace(dm_plan_t* P, unsigned long long* trace_dev, int cluster, int step) {
  if (!P) return fail(DM_ERR_ARG, "null argument");
  P->trace_dev = trace_dev; P->trace_cluster = cluster; P->trace_step = step;
  return DM_OK;
}

extern "C" int dm_plan_last_launch_count(const dm_plan_t* P) { return P ? P->last_launches : 0; }

extern "C" int dm_plan_set_sample_offset(dm_plan_t* P, long long first_sample) {
  if (!P || first_sample < 0) return fail(DM_ERR_ARG, "bad sample offset");
  if (((long long)P->H * P->T) % 4) return fail(DM_ERR_ARG, "H*T must be a multiple of 4");
  P->sample_offset = first_sample;
  return DM_OK;
}
