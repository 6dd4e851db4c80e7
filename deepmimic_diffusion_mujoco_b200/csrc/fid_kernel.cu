// K6 on the tensor cores: motion-FID moments  sum[D] += sum_n x_n,  sxx[D,D] += X^T X  (metrics/fid_score.py:24-43
// restated as streaming sums).
//
// X is fp32 [N, D] with the SAMPLE index as the contraction dimension, i.e. both GEMM operands are "MN-major" in
// memory.  A pre-pass transposes a chunk of samples into K-major fp16 hi/lo planes Xt[plane][D_pad][N_pad] (one read of
// X, one write of the same size; it also produces the column sums), after which the SYRK is a plain K-major
// 128 x 128 x 64 tcgen05 pipeline: TMA (SWIZZLE_128B) -> 3-stage mbarrier ring -> three kind::f16 MMAs per k-step
// (hi*hi + hi*lo + lo*hi, fp32 accumulate in TMEM: ~2^-22 relative per product).  Only upper-triangular tiles are
// computed.  Each CTA accumulates at most 4096 samples in fp32; the per-split partial tiles are combined in fp64 by
// a small reduction kernel that also mirrors the off-diagonal tiles.
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/dm_b200.h"
#include "ptx.cuh"

namespace dm {

int fail(int code, const char* fmt, ...);
int check_device();
int encode_map(CUtensorMap* m, void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
               const uint32_t* box, int swz);

constexpr int kFidT = 128;         // output tile
constexpr int kFidKc = 64;         // samples per pipeline stage: 64 fp16 = one 128-byte swizzled row
constexpr int kFidStages = 3;
constexpr int kFidTileBytes = kFidT * kFidKc * 2;           // 16 KB
constexpr int kFidStageBytes = 4 * kFidTileBytes;           // A hi, A lo, B hi, B lo
constexpr int kFidThreads = 192;   // warps 0-3: epilogue (TMEM lane quarters), warp 4: TMA producer, warp 5: MMA issuer
constexpr int kFidChunk = 8192;    // samples transposed per pass (bounds the scratch at ~75 MB for D = 2240)

struct FidBars {
  uint64_t full[kFidStages];
  uint64_t empty[kFidStages];
  uint64_t acc_full;
  uint32_t tmem_slot;
};

// ---- pre-pass: transpose + fp16 hi/lo split + column sums --------------------------------------------------
// block (64 columns d) x (64 samples k); 256 threads
__global__ void __launch_bounds__(256) fid_transpose_kernel(const float* __restrict__ X, int n0, int nc, int D, int D_pad,
                                                            int N_pad, __half* __restrict__ xt, double* __restrict__ sum) {
  __shared__ float tile[64][65];
  __shared__ float csum[4][64];
  const int d0 = blockIdx.x * 64, k0 = blockIdx.y * 64;
  const int dd = threadIdx.x & 63, kq = threadIdx.x >> 6;
  float cs = 0.f;
#pragma unroll 4
  for (int k = kq; k < 64; k += 4) {
    float v = 0.f;
    if (k0 + k < nc && d0 + dd < D) v = __ldcs(X + (size_t)(n0 + k0 + k) * D + d0 + dd);
    tile[k][dd] = v;
    cs += v;
  }
  csum[kq][dd] = cs;
  __syncthreads();
  if (threadIdx.x < 64 && d0 + threadIdx.x < D)
    atomicAdd(sum + d0 + threadIdx.x, (double)csum[0][threadIdx.x] + (double)csum[1][threadIdx.x] +
                                          (double)csum[2][threadIdx.x] + (double)csum[3][threadIdx.x]);
  // write: row d, 64 consecutive k as 32 half2
  const size_t plane = (size_t)D_pad * N_pad;
  const int kp = (threadIdx.x & 31) * 2, dr = threadIdx.x >> 5;
#pragma unroll 2
  for (int d = dr; d < 64; d += 8) {
    __half2 hi, lo;
    split_half2(tile[kp][d], tile[kp + 1][d], hi, lo);
    const size_t off = (size_t)(d0 + d) * N_pad + k0 + kp;
    *reinterpret_cast<__half2*>(xt + off) = hi;
    *reinterpret_cast<__half2*>(xt + plane + off) = lo;
  }
}

// ---- SYRK tile kernel ------------------------------------------------------------------------------------------
// grid (upper-triangular tile index, split); partial[split][tile][128][128] fp32
__global__ void __launch_bounds__(kFidThreads) fid_syrk_kernel(const __grid_constant__ CUtensorMap xt_map, int nt,
                                                               int chunks_total, int chunks_per_split,
                                                               float* __restrict__ partial) {
  extern __shared__ __align__(1024) uint8_t fid_smem_raw[];
  // 1024-byte alignment of the operand stages (SWIZZLE_128B atoms)
  uint8_t* stages = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(fid_smem_raw) + 1023) & ~uintptr_t(1023));
  FidBars& sh = *reinterpret_cast<FidBars*>(stages + kFidStages * kFidStageBytes);
  const int warp = warp_idx_sync();
  // upper-triangular tile (ti <= tj) from the linear index
  int ti = 0, rem = blockIdx.x;
  while (rem >= nt - ti) { rem -= nt - ti; ++ti; }
  const int tj = ti + rem;
  const int c0 = blockIdx.y * chunks_per_split;
  const int c1 = min(chunks_total, c0 + chunks_per_split);
  const int n_iter = c1 - c0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kFidStages; ++s) { mbar_init(&sh.full[s], 1); mbar_init(&sh.empty[s], 1); }
    mbar_init(&sh.acc_full, 1);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc(&sh.tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh.tmem_slot;

  if (warp == 4) {
    // ---------------- TMA producer ----------------
    if (elect_one_sync()) {
      tma_prefetch_desc(&xt_map);
      for (int it = 0; it < n_iter; ++it) {
        const int s = it % kFidStages, r = it / kFidStages;
        mbar_wait(&sh.empty[s], (r & 1) ^ 1, 0x601);
        mbar_arrive_expect_tx(&sh.full[s], kFidStageBytes);
        const uint32_t dst = smem_u32(stages + s * kFidStageBytes);
        const int k0 = (c0 + it) * kFidKc;
        tma_load_4d(dst, &xt_map, &sh.full[s], k0, ti * kFidT, 0, 0);
        tma_load_4d(dst + kFidTileBytes, &xt_map, &sh.full[s], k0, ti * kFidT, 1, 0);
        tma_load_4d(dst + 2 * kFidTileBytes, &xt_map, &sh.full[s], k0, tj * kFidT, 0, 0);
        tma_load_4d(dst + 3 * kFidTileBytes, &xt_map, &sh.full[s], k0, tj * kFidT, 1, 0);
      }
    }
  } else if (warp == 5) {
    // ---------------- MMA issuer ----------------
    const uint32_t idesc = umma_idesc_f16(kFidT, kFidT);
    for (int it = 0; it < n_iter; ++it) {
      const int s = it % kFidStages, r = it / kFidStages;
      mbar_wait(&sh.full[s], r & 1, 0x602);
      tc_fence_after();
      if (elect_one_sync()) {
        const uint32_t a_hi = smem_u32(stages + s * kFidStageBytes), a_lo = a_hi + kFidTileBytes;
        const uint32_t b_hi = a_hi + 2 * kFidTileBytes, b_lo = a_hi + 3 * kFidTileBytes;
#pragma unroll
        for (int ks = 0; ks < kFidKc / 16; ++ks) {
          const uint32_t o = 32u * ks;   // 16 fp16 along K inside the 128-byte swizzled row
          const uint64_t dah = umma_smem_desc(a_hi + o, 1024, 2), dal = umma_smem_desc(a_lo + o, 1024, 2);
          const uint64_t dbh = umma_smem_desc(b_hi + o, 1024, 2), dbl = umma_smem_desc(b_lo + o, 1024, 2);
          umma_f16(tmem_base, dal, dbh, idesc, !(it == 0 && ks == 0));
          umma_f16(tmem_base, dah, dbl, idesc, 1);
          umma_f16(tmem_base, dah, dbh, idesc, 1);
        }
        umma_commit(&sh.empty[s]);
        if (it == n_iter - 1) umma_commit(&sh.acc_full);
      }
      __syncwarp();
    }
  } else {
    // ---------------- epilogue: TMEM -> fp32 partial tile ----------------
    mbar_wait(&sh.acc_full, 0, 0x603);
    tc_fence_after();
    const int lane = threadIdx.x & 31;
    const int row = warp * 32 + lane;
    float* dst = partial + (((size_t)blockIdx.y * gridDim.x + blockIdx.x) * kFidT + row) * kFidT;
#pragma unroll 1
    for (int c = 0; c < kFidT / 32; ++c) {
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int q = 0; q < 8; ++q)
        __stcs(reinterpret_cast<float4*>(dst + c * 32) + q, make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc(tmem_base, 128);
}

// ---- combine the splits in fp64, add into sxx (and its mirror) -------------------------------------------------
__global__ void __launch_bounds__(256) fid_reduce_kernel(const float* __restrict__ partial, int n_tiles, int splits, int nt,
                                                         int D, double* __restrict__ sxx) {
  const int tile = blockIdx.x;
  int ti = 0, rem = tile;
  while (rem >= nt - ti) { rem -= nt - ti; ++ti; }
  const int tj = ti + rem;
  // blockIdx.y: one eighth (16 rows) of the tile, so that the grid covers the machine (171 tiles alone are 0.14 waves)
  const int e0 = blockIdx.y * (kFidT * kFidT / 8);
  for (int e = e0 + threadIdx.x; e < e0 + kFidT * kFidT / 8; e += 256) {
    const int r = e / kFidT, c = e % kFidT;
    const int i = ti * kFidT + r, j = tj * kFidT + c;
    if (i >= D || j >= D) continue;
    double a = 0.0;
    for (int s = 0; s < splits; ++s) a += (double)__ldcs(partial + ((size_t)s * n_tiles + tile) * (kFidT * kFidT) + e);
    sxx[(size_t)i * D + j] += a;
    if (ti != tj) sxx[(size_t)j * D + i] += a;
  }
}

}  // namespace dm

using namespace dm;

static cudaMemPool_t fid_pool(int dev) {
  // library-owned stream-ordered pool that never trims (the default pool frees at every synchronisation)
  static cudaMemPool_t pools[64] = {};
  if (dev < 0 || dev >= 64) return nullptr;
  if (!pools[dev]) {
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = dev;
    if (cudaMemPoolCreate(&pools[dev], &props) != cudaSuccess) { pools[dev] = nullptr; return nullptr; }
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
  }
  return pools[dev];
}

extern "C" int dm_fid_moments(const float* X_dev, int N, int D, double* sum_dev, double* sxx_dev, dm_stream_t st) {
  if (!X_dev || !sum_dev || !sxx_dev) return fail(DM_ERR_ARG, "null argument");
  if (N < 1 || D < 1) return fail(DM_ERR_ARG, "need N >= 1, D >= 1");
  int rc = check_device();
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  int dev = 0;
  cudaGetDevice(&dev);
  cudaMemPool_t pool = fid_pool(dev);
  if (!pool) return fail(DM_ERR_CUDA, "cudaMemPoolCreate failed");
  const int nt = (D + kFidT - 1) / kFidT;
  const int D_pad = nt * kFidT;
  const int n_tiles = nt * (nt + 1) / 2;
  const int max_nc = N < kFidChunk ? N : kFidChunk;
  const int N_pad_max = (max_nc + kFidKc - 1) / kFidKc * kFidKc;
  const int max_splits = 16;
  __half* xt = nullptr;
  float* partial = nullptr;
  cudaError_t e = cudaMallocFromPoolAsync(reinterpret_cast<void**>(&xt), (size_t)2 * D_pad * N_pad_max * sizeof(__half), pool, s);
  if (e == cudaSuccess)
    e = cudaMallocFromPoolAsync(reinterpret_cast<void**>(&partial), (size_t)max_splits * n_tiles * kFidT * kFidT * sizeof(float), pool, s);
  if (e != cudaSuccess) { if (xt) cudaFreeAsync(xt, s); return fail(DM_ERR_CUDA, "cudaMallocFromPoolAsync: %s", cudaGetErrorString(e)); }
  const size_t smem = (size_t)kFidStages * kFidStageBytes + sizeof(FidBars) + 1024;
  e = cudaFuncSetAttribute(fid_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int status = DM_OK;
  for (int n0 = 0; n0 < N && e == cudaSuccess && status == DM_OK; n0 += kFidChunk) {
    const int nc = N - n0 < kFidChunk ? N - n0 : kFidChunk;
    const int N_pad = (nc + kFidKc - 1) / kFidKc * kFidKc;
    dim3 tg(D_pad / 64, N_pad / 64);
    fid_transpose_kernel<<<tg, 256, 0, s>>>(X_dev, n0, nc, D, D_pad, N_pad, xt, sum_dev);
    // tensor map over this chunk's planes: (k, d, plane, 1), box 64 x 128 x 1 x 1, 128-byte swizzle
    CUtensorMap map;
    const uint64_t dims[4] = {(uint64_t)N_pad, (uint64_t)D_pad, 2, 1};
    const uint64_t strides[3] = {(uint64_t)N_pad * 2, (uint64_t)D_pad * N_pad * 2, (uint64_t)2 * D_pad * N_pad * 2};
    const uint32_t box[4] = {kFidKc, kFidT, 1, 1};
    status = encode_map(&map, xt, 4, dims, strides, box, 2);
    if (status != DM_OK) break;
    // splits: every CTA accumulates <= 4096 samples in fp32; pick the count that minimises waves x chunks per split
    const int chunks = N_pad / kFidKc;
    int best = 1;
    long long best_cost = -1;
    for (int sp = 1; sp <= max_splits && sp <= chunks; ++sp) {
      const int per = (chunks + sp - 1) / sp;
      if (per * kFidKc > 4096) continue;
      const int real = (chunks + per - 1) / per;
      const long long waves = ((long long)n_tiles * real + 147) / 148;
      const long long cost = waves * (per + 6);   // + fixed pipeline fill / epilogue per CTA, in chunk units
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = sp; }
    }
    const int per = (chunks + best - 1) / best;
    const int splits = (chunks + per - 1) / per;
    fid_syrk_kernel<<<dim3(n_tiles, splits), kFidThreads, smem, s>>>(map, nt, chunks, per, partial);
    fid_reduce_kernel<<<dim3(n_tiles, 8), 256, 0, s>>>(partial, n_tiles, splits, nt, D, sxx_dev);
    e = cudaGetLastError();
  }
  cudaFreeAsync(partial, s);
  cudaFreeAsync(xt, s);
  if (status != DM_OK) return status;
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "fid_moments launch: %s", cudaGetErrorString(e));
  return DM_OK;
}
