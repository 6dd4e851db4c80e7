// K5: batched humanoid forward kinematics + DeepMimic tracking reward; K6: motion-FID moment accumulation.
//
// FK follows MuJoCo mj_kinematics on /root/reference/diffusion/assets/dp_env_v2.xml:20-107 (the reference
// only reaches FK through mujoco_py sim.forward(), mocap_player.py:67-73): free root = qpos[0:7]; a child body
// starts at parent_pos + R_parent * body_offset with the parent's orientation; each hinge, in document order,
// rotates the frame about its body-frame axis through its anchor.
// Reward: SURVEY.md 8(a) a15 (this repo's spec; weights from utils/mocap_util.py:26-29).
// FID moments: metrics/fid_score.py:24-43 restated as streaming sums.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "../../include/dm_b200.h"
#include "fk_math.cuh"

namespace dm {

int fail(int code, const char* fmt, ...);
int check_device();

// ------------------------------------------------------------------------------------------------
// K5.  One thread per frame; the arithmetic lives in fk_math.cuh.  A block walks tiles of 128 consecutive frames
// (grid-stride, grid = a multiple of the SM count): the tile's qpos rows are staged into shared memory with
// coalesced 16-byte loads and read back row-wise (row stride 35 words: conflict-free); the reference clip's
// features sit in shared memory for the whole kernel (row stride 92 words).  Registers are capped at 128 so that
// 16 warps per SM hide the latency of the dependent FK chain.
// ------------------------------------------------------------------------------------------------
constexpr int kFkThreads = 128;
constexpr int kFkBlocksPerSm = 4;

__global__ void ref_features_kernel(const float* __restrict__ ref, int H, float inv_dt, float* __restrict__ feat) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= H) return;
  float* fo = feat + (size_t)h * fk::kFeatStride;
  for (int i = 86; i < fk::kFeatStride; ++i) fo[i] = 0.f;
  fk::frame_pass<true, false>(ref + (size_t)h * 35, ref + (size_t)(h > 0 ? h - 1 : 0) * 35, inv_dt, nullptr, fo, nullptr, nullptr);
}

template <bool kBody, bool kFeatSmem>
__global__ void __launch_bounds__(kFkThreads, kFkBlocksPerSm)
fk_reward_kernel(const float* __restrict__ qpos, const float* __restrict__ feat, long long n_frames, int H, float inv_dt,
                 int vec_ok, int body_vec_ok, float* __restrict__ body_pos, float* __restrict__ reward, double* __restrict__ sums) {
  extern __shared__ __align__(16) float fk_smem[];
  __shared__ float red[6][kFkThreads / 32];
  float* sq = fk_smem;                              // [128][35] rows of this tile
  float* sprev = fk_smem + kFkThreads * 35;         // the frame before the tile (35 floats, padded to 36)
  float* sbody = fk_smem + kFkThreads * 35 + 36;    // [128][45] when kBody: positions leave through coalesced stores
  float* sfeat = sbody + (kBody ? kFkThreads * fk::kBodies * 3 : 0);   // [H][92] when kFeatSmem
  if (kFeatSmem) {
    const float4* src = reinterpret_cast<const float4*>(feat);
    float4* dst = reinterpret_cast<float4*>(sfeat);
    for (int i = threadIdx.x; i < H * (fk::kFeatStride / 4); i += kFkThreads) dst[i] = __ldg(src + i);
  }
  const long long n_tiles = (n_frames + kFkThreads - 1) / kFkThreads;
  float s_r = 0.f, s_r2 = 0.f, s_p = 0.f, s_v = 0.f, s_e = 0.f, s_c = 0.f;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long f0 = tile * kFkThreads;
    const int nf = (int)min((long long)kFkThreads, n_frames - f0);
    __syncthreads();   // previous tile fully consumed (also orders the feature copy before first use)
    {
      const float* src = qpos + f0 * 35;
      const int n = nf * 35;
      int done = 0;
      if (vec_ok) {   // f0 * 140 bytes is a multiple of 16: the tile starts 16-byte aligned when qpos does
        const int n4 = n >> 2;
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(sq);
        for (int i = threadIdx.x; i < n4; i += kFkThreads) d4[i] = __ldcs(s4 + i);
        done = n4 << 2;
      }
      for (int i = done + threadIdx.x; i < n; i += kFkThreads) sq[i] = __ldcs(src + i);
      if (f0 > 0 && threadIdx.x < 35) sprev[threadIdx.x] = __ldg(src - 35 + threadIdx.x);
    }
    __syncthreads();
    if ((int)threadIdx.x < nf) {
      const long long fi = f0 + threadIdx.x;
      const int h = (int)(((unsigned)(f0 % H) + threadIdx.x) % (unsigned)H);   // one 64-bit modulo per tile (uniform), 32-bit per thread
      const float* q = sq + threadIdx.x * 35;
      const float* qp = h == 0 ? q : (threadIdx.x == 0 ? sprev : q - 35);
      const float* g = (kFeatSmem ? sfeat : feat) + (size_t)h * fk::kFeatStride;
      float acc[4];
      fk::frame_pass<false, kBody>(q, qp, inv_dt, g, nullptr, kBody ? sbody + threadIdx.x * (fk::kBodies * 3) : nullptr, acc);
      const float tp = fk::fast_exp(-2.f * acc[0]), tv = fk::fast_exp(-0.1f * acc[1]);
      const float te = fk::fast_exp(-40.f * acc[2]), tc = fk::fast_exp(-10.f * acc[3]);
      const float r = 0.65f * tp + 0.1f * tv + 0.15f * te + 0.1f * tc;
      __stcs(reward + fi, r);
      s_r += r; s_r2 = fmaf(r, r, s_r2); s_p += tp; s_v += tv; s_e += te; s_c += tc;
    }
    if (kBody) {
      __syncthreads();
      float* dst = body_pos + (size_t)f0 * (fk::kBodies * 3);
      const int n = nf * fk::kBodies * 3;
      int done = 0;
      if (body_vec_ok) {   // f0 * 180 bytes is a multiple of 16
        const int n4 = n >> 2;
        for (int i = threadIdx.x; i < n4; i += kFkThreads) __stcs(reinterpret_cast<float4*>(dst) + i, reinterpret_cast<const float4*>(sbody)[i]);
        done = n4 << 2;
      }
      for (int i = done + threadIdx.x; i < n; i += kFkThreads) __stcs(dst + i, sbody[i]);
    }
  }
  // block reduction -> one double atomic per statistic per block
  float vals[6] = {s_r, s_r2, s_p, s_v, s_e, s_c};
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float v = vals[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    double s = 0.0;
    for (int w = 0; w < kFkThreads / 32; ++w) s += (double)red[threadIdx.x][w];
    atomicAdd(&sums[threadIdx.x], s);
  }
  if (threadIdx.x == 6 && blockIdx.x == 0) atomicAdd(&sums[6], (double)n_frames);
}

// ------------------------------------------------------------------------------------------------
// Batched window statistics for SiFID (eval.py:59-77: per generated sample, mean and covariance over its W windows
// of D = window_size * n_feats values, fid_score.py:24-43 arithmetic in fp32).  One CTA per (sample, 8-row band of
// the covariance); the sample's W x D window matrix is centred in shared memory.  HBM-bound on the D x D output.
// ------------------------------------------------------------------------------------------------
constexpr int kWinRows = 32;   // covariance rows per CTA
constexpr int kWinMaxW = 32;   // windows per sample held in registers (H <= 320 at window 10, step 10)

template <int kWMax>   // windows held in registers: the smallest of 8 / 16 / 32 that covers W (predicated-off FMAs still issue)
__global__ void __launch_bounds__(256) window_moments_kernel(const float* __restrict__ X, int H, int F, int W, int win,
                                                             int step, float* __restrict__ mu, float* __restrict__ sigma) {
  extern __shared__ float wsm[];   // [W][D] centred windows
  const int D = win * F;
  const int n = blockIdx.y;
  const float* xs = X + (size_t)n * H * F;
  // window w covers frames [w*step, w*step + win): D contiguous floats starting at w*step*F
  for (int w = 0; w < W; ++w)
    for (int d = threadIdx.x; d < D; d += blockDim.x) wsm[w * D + d] = __ldg(xs + (size_t)w * step * F + d);
  __syncthreads();
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float m = 0.f;
    for (int w = 0; w < W; ++w) m += wsm[w * D + d];
    m /= (float)W;
    for (int w = 0; w < W; ++w) wsm[w * D + d] -= m;
    if (blockIdx.x == 0) mu[(size_t)n * D + d] = m;
  }
  __syncthreads();
  const float inv = 1.f / (float)(W - 1);
  const int i0 = blockIdx.x * kWinRows;
  const int rows = min(kWinRows, D - i0);
  // thread <-> column j: its W centred values stay in registers; the row values are shared-memory broadcasts, so an
  // output element costs W FMAs + W broadcast loads and the D x D block leaves in coalesced streaming stores
  for (int j = threadIdx.x; j < D; j += blockDim.x) {
    float bcol[kWMax];
#pragma unroll
    for (int w = 0; w < kWMax; ++w) bcol[w] = w < W ? wsm[w * D + j] : 0.f;
    float* out = sigma + ((size_t)n * D + i0) * D + j;
    for (int r = 0; r < rows; ++r) {
      const float* arow = wsm + i0 + r;
      float acc = 0.f;
#pragma unroll
      for (int w = 0; w < kWMax; ++w)
        if (w < W) acc = fmaf(arow[w * D], bcol[w], acc);
      __stcs(out + (size_t)r * D, acc * inv);
    }
  }
}

}  // namespace dm

using namespace dm;

extern "C" int dm_fk_reward(const float* qpos_dev, const float* ref_dev, int N, int H, float dt, float* body_pos_dev,
                            float* reward_dev, double* sums_dev, dm_stream_t st) {
  if (!qpos_dev || !ref_dev || !reward_dev || !sums_dev) return fail(DM_ERR_ARG, "null argument");
  if (N < 1 || H < 1 || !(dt > 0.f)) return fail(DM_ERR_ARG, "need N >= 1, H >= 1, dt > 0");
  int rc = check_device();
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  // Scratch for the reference clip's features comes from a library-owned stream-ordered pool that never trims:
  // the default pool returns its memory to the driver at every synchronisation, which made each call pay a real
  // allocation (~240 us measured on B200, 6x the kernel itself at 8192 samples).
  static cudaMemPool_t pools[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return fail(DM_ERR_ARG, "device ordinal %d out of range", dev);
  if (!pools[dev]) {
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = dev;
    cudaError_t pe = cudaMemPoolCreate(&pools[dev], &props);
    if (pe != cudaSuccess) { pools[dev] = nullptr; return fail(DM_ERR_CUDA, "cudaMemPoolCreate: %s", cudaGetErrorString(pe)); }
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
  }
  float* feat = nullptr;
  cudaError_t e = cudaMallocFromPoolAsync(reinterpret_cast<void**>(&feat), (size_t)H * fk::kFeatStride * sizeof(float), pools[dev], s);
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "cudaMallocFromPoolAsync: %s", cudaGetErrorString(e));
  ref_features_kernel<<<(H + 31) / 32, 32, 0, s>>>(ref_dev, H, 1.f / dt, feat);
  const long long nfr = (long long)N * H;
  const long long tiles = (nfr + kFkThreads - 1) / kFkThreads;
  const bool body = body_pos_dev != nullptr;
  const size_t base_smem = (size_t)(kFkThreads * 35 + 36 + (body ? kFkThreads * fk::kBodies * 3 : 0)) * sizeof(float);
  const size_t feat_smem = (size_t)H * fk::kFeatStride * sizeof(float);
  const bool feat_in_smem = base_smem + feat_smem <= 200 * 1024;
  const size_t smem = base_smem + (feat_in_smem ? feat_smem : 0);
  int blocks_per_sm = kFkBlocksPerSm;
  while (blocks_per_sm > 1 && (smem + 1024) * blocks_per_sm > 227 * 1024) --blocks_per_sm;
  const unsigned grid = (unsigned)(tiles < 148LL * blocks_per_sm ? tiles : 148LL * blocks_per_sm);
  const int vec_ok = (reinterpret_cast<uintptr_t>(qpos_dev) & 15) == 0;
  const int body_vec_ok = (reinterpret_cast<uintptr_t>(body_pos_dev) & 15) == 0;
  auto launch = [&](auto kern) -> cudaError_t {
    if (smem > 48 * 1024) {
      cudaError_t le = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (le != cudaSuccess) return le;
    }
    kern<<<grid, kFkThreads, smem, s>>>(qpos_dev, feat, nfr, H, 1.f / dt, vec_ok, body_vec_ok, body_pos_dev, reward_dev, sums_dev);
    return cudaGetLastError();
  };
  if (body) e = feat_in_smem ? launch(fk_reward_kernel<true, true>) : launch(fk_reward_kernel<true, false>);
  else e = feat_in_smem ? launch(fk_reward_kernel<false, true>) : launch(fk_reward_kernel<false, false>);
  cudaFreeAsync(feat, s);
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "fk_reward launch: %s", cudaGetErrorString(e));
  return DM_OK;
}

// ------------------------------------------------------------------------------------------------
// dm_fk_reward_noalloc: the same kernels as dm_fk_reward with the reference-feature scratch in a file-scope device
// array instead of a pool allocation -- for callers that must not allocate (CUDA graph capture, serving loops).
// No counterpart in the reference.
//   * kNoAllocMaxH frames of kFeatStride floats = 376 832 bytes, held by every GPU that loads the library until exit.
//   * In HOST code `g_fk_noalloc_feat` is only a stand-in for the symbol: the device address comes from
//     cudaGetSymbolAddress, fetched once per device with that device current (never pass the name to a kernel).
//   * Every call shares the array and it keeps the previous call's contents.  Nothing has to start at zero:
//     ref_features_kernel rewrites all kFeatStride floats of the H rows a call uses (padding included) before
//     fk_reward_kernel reads them, in stream order.  Calls that overlap on the GPU (different streams, or graph
//     branches without a dependency) race on the array: serialise them.
// ------------------------------------------------------------------------------------------------
constexpr int kNoAllocMaxH = DM_FK_NOALLOC_MAX_H;
static __device__ __align__(256) float g_fk_noalloc_feat[kNoAllocMaxH * dm::fk::kFeatStride];

static int fk_noalloc_scratch(float** out) {
  static std::mutex mu;
  static float* addr[64] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= 64) return fail(DM_ERR_ARG, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_fk_noalloc_feat);   // this device's copy of the array (device `dev` is current)
    if (e != cudaSuccess || !p) return fail(DM_ERR_CUDA, "cudaGetSymbolAddress(g_fk_noalloc_feat): %s", cudaGetErrorString(e));
    addr[dev] = static_cast<float*>(p);
  }
  *out = addr[dev];
  return DM_OK;
}

extern "C" int dm_fk_reward_noalloc(const float* qpos_dev, const float* ref_dev, int N, int H, float dt, float* body_pos_dev,
                                    float* reward_dev, double* sums_dev, dm_stream_t st) {
  if (!qpos_dev || !ref_dev || !reward_dev || !sums_dev) return fail(DM_ERR_ARG, "null argument");
  if (N < 1 || H < 1 || !(dt > 0.f)) return fail(DM_ERR_ARG, "need N >= 1, H >= 1, dt > 0");
  if (H > kNoAllocMaxH)
    return fail(DM_ERR_ARG, "dm_fk_reward_noalloc: horizon %d exceeds the static scratch (DM_FK_NOALLOC_MAX_H = %d frames); "
                            "use dm_fk_reward", H, kNoAllocMaxH);
  int rc = check_device();
  if (rc) return rc;
  float* feat = nullptr;
  if ((rc = fk_noalloc_scratch(&feat))) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  ref_features_kernel<<<(H + 31) / 32, 32, 0, s>>>(ref_dev, H, 1.f / dt, feat);
  const long long nfr = (long long)N * H;
  const long long tiles = (nfr + kFkThreads - 1) / kFkThreads;
  const bool body = body_pos_dev != nullptr;
  const size_t base_smem = (size_t)(kFkThreads * 35 + 36 + (body ? kFkThreads * fk::kBodies * 3 : 0)) * sizeof(float);
  const size_t feat_smem = (size_t)H * fk::kFeatStride * sizeof(float);
  const bool feat_in_smem = base_smem + feat_smem <= 200 * 1024;
  const size_t smem = base_smem + (feat_in_smem ? feat_smem : 0);
  int blocks_per_sm = kFkBlocksPerSm;
  while (blocks_per_sm > 1 && (smem + 1024) * blocks_per_sm > 227 * 1024) --blocks_per_sm;
  const unsigned grid = (unsigned)(tiles < 148LL * blocks_per_sm ? tiles : 148LL * blocks_per_sm);
  const int vec_ok = (reinterpret_cast<uintptr_t>(qpos_dev) & 15) == 0;
  const int body_vec_ok = (reinterpret_cast<uintptr_t>(body_pos_dev) & 15) == 0;
  auto launch = [&](auto kern) -> cudaError_t {
    if (smem > 48 * 1024) {
      cudaError_t le = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (le != cudaSuccess) return le;
    }
    kern<<<grid, kFkThreads, smem, s>>>(qpos_dev, feat, nfr, H, 1.f / dt, vec_ok, body_vec_ok, body_pos_dev, reward_dev, sums_dev);
    return cudaGetLastError();
  };
  cudaError_t e;
  if (body) e = feat_in_smem ? launch(fk_reward_kernel<true, true>) : launch(fk_reward_kernel<true, false>);
  else e = feat_in_smem ? launch(fk_reward_kernel<false, true>) : launch(fk_reward_kernel<false, false>);
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "fk_reward launch: %s", cudaGetErrorString(e));
  return DM_OK;
}

extern "C" int dm_window_moments(const float* X_dev, int N, int H, int F, int window, int step, float* mu_dev, float* sigma_dev,
                                 dm_stream_t st) {
  if (!X_dev || !mu_dev || !sigma_dev) return fail(DM_ERR_ARG, "null argument");
  if (N < 1 || H < 1 || F < 1 || window < 1 || step < 1 || window > H) return fail(DM_ERR_ARG, "need N, F, step >= 1 and 1 <= window <= H");
  const int W = (H - window) / step + 1;   // offsets 0, step, 2 step, ... <= H - window (eval.py:15-21)
  if (W < 2) return fail(DM_ERR_ARG, "need at least two windows per sample (covariance divides by W - 1)");
  if (W > kWinMaxW) return fail(DM_ERR_ARG, "%d windows per sample; at most %d are supported", W, kWinMaxW);
  int rc = check_device();
  if (rc) return rc;
  const int D = window * F;
  const size_t smem = (size_t)W * D * sizeof(float);
  if (smem > 200 * 1024) return fail(DM_ERR_ARG, "window matrix %d x %d does not fit in shared memory", W, D);
  if (N > 65535) return fail(DM_ERR_ARG, "at most 65535 samples per call");
  dim3 grid((D + kWinRows - 1) / kWinRows, N);
  auto kern = W <= 8 ? window_moments_kernel<8> : (W <= 16 ? window_moments_kernel<16> : window_moments_kernel<32>);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail(DM_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  }
  kern<<<grid, 256, smem, static_cast<cudaStream_t>(st)>>>(X_dev, H, F, W, window, step, mu_dev, sigma_dev);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "window_moments launch: %s", cudaGetErrorString(e));
  return DM_OK;
}
