// K1 + K2: Conv1d-family implicit GEMM on tcgen05 (fp16 hi/lo split operands, fp32 accumulate in TMEM), TMA
// fed, with the reference block's epilogue fused: bias -> GroupNorm(8) -> Mish -> + time bias -> + residual.
//
// Data layout: activations are fp16 planes [plane][G][L][16][C] (plane 0 = hi, 1 = lo,
// value = hi + lo).  A CTA owns S (8 or 16) samples of one group for ALL L positions, so its GEMM rows are
// ordered (l, b): row = l * S + b.  A conv tap that reads position l + d is then the SAME shared-memory tile
// viewed d * S rows further down -- a whole number of 8-row swizzle atoms -- so the five taps of a k=5 conv
// are five UMMA descriptors over one TMA-loaded box (halo rows zero-filled by TMA out-of-bounds handling).
// The whole per-sample conv output lives in TMEM until GroupNorm statistics are known; no cross-CTA reduce.
//
// Reference: Conv1dBlock helpers.py:50-69; ResidualTemporalBlock temporal.py:37-46; Downsample1d/Upsample1d
// helpers.py:32-47 (under /root/reference/diffusion/diffuser/models/).
#include "conv_job.h"
#include "ptx.cuh"

namespace dm {

__device__ __forceinline__ float mish_f(float x) {
  // x * tanh(softplus(x)) == x * n(n+2) / (n(n+2) + 2), n = e^x ; torch uses softplus threshold 20
  if (x > 20.f) return x;
  float n = __expf(x);
  float a = n * (n + 2.f);
  return x * __fdividef(a, a + 2.f);
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__global__ void __launch_bounds__(kConvThreads, 2) conv_gemm_kernel(const __grid_constant__ ConvJob job) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_smem = smem;
  uint8_t* w_smem = smem + (size_t)job.n_a_stages * job.a_stage_bytes;

  __shared__ uint64_t a_full[4], a_empty[4], w_full[8], w_empty[8], acc_full;
  __shared__ uint32_t tmem_slot;
  __shared__ float s_par[4][256];      // bias, gamma, beta, time-bias for this CTA's Nc channels
  __shared__ float s_stat[16][8][2];   // per (sample, group): sum, sum of squares
  __shared__ float s_part[4][16][8][2]; // per-warp partials, combined in a fixed order (deterministic)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int S = job.S;
  const int subs = kGroupSamples / S;
  const int g = blockIdx.x / subs;
  const int b0 = (blockIdx.x % subs) * S;
  const int n0 = blockIdx.y * job.Nc;
  const int Nc = job.Nc;
  const int MT = job.n_ltiles * job.Mt;

  if (threadIdx.x == 0) {
    for (int i = 0; i < job.n_a_stages; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < job.n_w_stages; ++i) { mbar_init(&w_full[i], 1); mbar_init(&w_empty[i], 1); }
    mbar_init(&acc_full, 1);
    fence_barrier_init();
  }
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&job.a_map[0]);
    if (job.nsrc > 1) tma_prefetch_desc(&job.a_map[1]);
    tma_prefetch_desc(&job.w_map);
  }
  if (warp == 5) tmem_alloc(&tmem_slot, job.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 4) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      int as = 0, aph = 0, ws = 0, wph = 0;
      for (int lt = 0; lt < job.n_ltiles; ++lt) {
        const int lstart = job.l_stride * (lt * job.Lt) - job.halo_lo;
        int cin_off = 0;
        for (int src = 0; src < job.nsrc; ++src) {
          for (int ch = 0; ch < job.src_chunks[src]; ++ch) {
            mbar_wait(&a_empty[as], aph ^ 1, 0x100 + as);
            mbar_arrive_expect_tx(&a_full[as], job.a_tx_bytes);
            tma_load_5d(a_smem + (size_t)as * job.a_stage_bytes, &job.a_map[src], &a_full[as], ch * job.Cc, b0,
                        lstart, g, 0);
            for (int t = 0; t < job.ntaps; ++t) {
              mbar_wait(&w_empty[ws], wph ^ 1, 0x200 + ws);
              mbar_arrive_expect_tx(&w_full[ws], job.w_tx_bytes);
              tma_load_4d(w_smem + (size_t)ws * job.w_stage_bytes, &job.w_map, &w_full[ws],
                          cin_off + ch * job.Cc, n0, job.taps[t].wtap, 0);
              if (++ws == job.n_w_stages) { ws = 0; wph ^= 1; }
            }
            if (++as == job.n_a_stages) { as = 0; aph ^= 1; }
          }
          cin_off += job.src_chunks[src] * job.Cc;
        }
      }
    }
  } else if (warp == 5) {
    // ================================ MMA issuer ==================================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_f16(128, Nc);
      const uint32_t sbo_a = job.l_stride * 8 * job.row_bytes;
      const uint32_t sbo_w = 8 * job.row_bytes;
      const int ksteps = job.Cc / 16;
      const int rows_per_mt = 128 / S;  // M positions per 128-row tile
      int as = 0, aph = 0, ws = 0, wph = 0;
      for (int lt = 0; lt < job.n_ltiles; ++lt) {
        int cidx = 0;
        for (int src = 0; src < job.nsrc; ++src) {
          for (int ch = 0; ch < job.src_chunks[src]; ++ch, ++cidx) {
            mbar_wait(&a_full[as], aph, 0x300 + as);
            tc_fence_after();
            const uint32_t a_base = smem_u32(a_smem + (size_t)as * job.a_stage_bytes);
            for (int t = 0; t < job.ntaps; ++t) {
              mbar_wait(&w_full[ws], wph, 0x400 + ws);
              tc_fence_after();
              const uint32_t w_base = smem_u32(w_smem + (size_t)ws * job.w_stage_bytes);
              const ConvTap tap = job.taps[t];
              for (int mt = 0; mt < job.Mt; ++mt) {
                const uint32_t d = tmem_base + (uint32_t)((tap.phase * MT + lt * job.Mt + mt) * Nc);
                const uint32_t a_row = (uint32_t)((job.l_stride * mt * rows_per_mt + tap.shift) * S);
                const uint32_t a_addr = a_base + a_row * job.row_bytes;
                for (int k = 0; k < ksteps; ++k) {
                  const uint32_t acc = !(cidx == 0 && tap.first && k == 0);
                  const uint64_t a_hi = umma_smem_desc(a_addr + k * 32, sbo_a, job.swz);
                  const uint64_t a_lo = umma_smem_desc(a_addr + job.a_plane_bytes + k * 32, sbo_a, job.swz);
                  const uint64_t w_hi = umma_smem_desc(w_base + k * 32, sbo_w, job.swz);
                  const uint64_t w_lo = umma_smem_desc(w_base + job.w_plane_bytes + k * 32, sbo_w, job.swz);
                  umma_f16(d, a_lo, w_hi, idesc, acc);
                  umma_f16(d, a_hi, w_lo, idesc, 1);
                  umma_f16(d, a_hi, w_hi, idesc, 1);
                }
              }
              umma_commit(&w_empty[ws]);
              if (++ws == job.n_w_stages) { ws = 0; wph ^= 1; }
            }
            umma_commit(&a_empty[as]);
            if (++as == job.n_a_stages) { as = 0; aph ^= 1; }
          }
        }
      }
      umma_commit(&acc_full);
    }
  } else {
    // ================================ epilogue (warps 0-3) =========================
    const int tid = threadIdx.x;       // 0..127 == TMEM lane
    const int b = tid & (S - 1);
    const bool gn = job.epi == EPI_GN_MISH;
    for (int i = tid; i < Nc; i += 128) {
      s_par[0][i] = job.bias[n0 + i];
      s_par[1][i] = gn ? job.gamma[n0 + i] : 1.f;
      s_par[2][i] = gn ? job.beta[n0 + i] : 0.f;
      s_par[3][i] = (job.tbias && job.tbias_stride == 0) ? job.tbias[n0 + i] : 0.f;
    }
    for (int i = tid; i < 4 * 16 * 8 * 2; i += 128) (&s_part[0][0][0][0])[i] = 0.f;
    named_bar_sync(1, 128);

    mbar_wait(&acc_full, 0, 0x500);
    tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
    const int nchunks = Nc >> 5;
    float v[32];

    if (gn) {
      // ---- pass 1: GroupNorm statistics over (channels of group) x (all L positions), per sample ----
      for (int cc = 0; cc < nchunks; ++cc) {
        float s[8], q[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) { s[r] = 0.f; q[r] = 0.f; }
        for (int T = 0; T < MT; ++T) {
          const int j = (T * 128 + tid) / S;
          tmem_ld32(trow + (uint32_t)(T * Nc + cc * 32), v);
          tmem_ld_wait();
          if (j < job.Lj) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float x = v[r * 4 + e] + s_par[0][cc * 32 + r * 4 + e];
                s[r] += x;
                q[r] = fmaf(x, x, q[r]);
              }
            }
          }
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          s[r] += __shfl_xor_sync(0xffffffffu, s[r], 16);
          q[r] += __shfl_xor_sync(0xffffffffu, q[r], 16);
          if (S == 8) {
            s[r] += __shfl_xor_sync(0xffffffffu, s[r], 8);
            q[r] += __shfl_xor_sync(0xffffffffu, q[r], 8);
          }
          if (lane < S) {
            // runs of 4 channels that fall in the same group are added in increasing r: fixed order
            const int gi = (cc * 32 + r * 4) / job.cg;
            s_part[warp][lane][gi][0] += s[r];
            s_part[warp][lane][gi][1] += q[r];
          }
        }
      }
      named_bar_sync(1, 128);
      for (int i = tid; i < 16 * 8 * 2; i += 128) {
        const float* pp = &s_part[0][0][0][0];
        (&s_stat[0][0][0])[i] = ((pp[i] + pp[256 + i]) + pp[512 + i]) + pp[768 + i];
      }
      named_bar_sync(1, 128);
    }

    // ---- pass 2: normalise, Mish, + time bias, + residual, split to fp16 hi/lo, store ----
    const float inv_n = 1.f / (float)(job.cg * job.Lj);
    const int bg = g * kGroupSamples + b0 + b;
    for (int T = 0; T < MT; ++T) {
      const int j = (T * 128 + tid) / S;
      const bool valid = j < job.Lj;
      for (int ph = 0; ph < job.nphase; ++ph) {
        const int lrow = job.nphase == 2 ? 2 * j + ph : j;
        const size_t rowoff = ((size_t)g * job.Lout + lrow) * kGroupSamples + b0 + b;
        for (int cc = 0; cc < nchunks; ++cc) {
          tmem_ld32(trow + (uint32_t)((ph * MT + T) * Nc + cc * 32), v);
          tmem_ld_wait();
          if (!valid) continue;
          const int cb = cc * 32;
#pragma unroll
          for (int r = 0; r < 8; ++r) {
            float mean = 0.f, rstd = 1.f;
            if (gn) {
              const int gi = (cb + r * 4) / job.cg;
              mean = s_stat[b][gi][0] * inv_n;
              const float var = fmaxf(s_stat[b][gi][1] * inv_n - mean * mean, 0.f);
              rstd = rsqrtf(var + 1e-5f);
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int c = cb + r * 4 + e;
              float x = v[r * 4 + e] + s_par[0][c];
              if (gn) {
                x = (x - mean) * rstd * s_par[1][c] + s_par[2][c];
                x = mish_f(x);
                x += s_par[3][c];
              }
              v[r * 4 + e] = x;
            }
          }
          if (gn && job.tbias && job.tbias_stride != 0) {
            const float* tb = job.tbias + (size_t)bg * job.tbias_stride + n0 + cb;
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] += __ldg(tb + e);
          }
          if (job.res) {
            const uint4* rh = reinterpret_cast<const uint4*>(job.res + rowoff * job.res_C + n0 + cb);
            const uint4* rl = reinterpret_cast<const uint4*>(job.res + job.res_plane_stride + rowoff * job.res_C + n0 + cb);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint4 hv = __ldg(rh + u), lv = __ldg(rl + u);
              const __half2* hh = reinterpret_cast<const __half2*>(&hv);
              const __half2* ll = reinterpret_cast<const __half2*>(&lv);
#pragma unroll
              for (int w = 0; w < 4; ++w) {
                const float2 fh = __half22float2(hh[w]), fl = __half22float2(ll[w]);
                v[u * 8 + w * 2] += fh.x + fl.x;
                v[u * 8 + w * 2 + 1] += fh.y + fl.y;
              }
            }
          }
          if (job.out) {
            uint4* oh = reinterpret_cast<uint4*>(job.out + rowoff * job.out_C + n0 + cb);
            uint4* ol = reinterpret_cast<uint4*>(job.out + job.out_plane_stride + rowoff * job.out_C + n0 + cb);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              uint4 hv, lv;
              __half2* hh = reinterpret_cast<__half2*>(&hv);
              __half2* ll = reinterpret_cast<__half2*>(&lv);
#pragma unroll
              for (int w = 0; w < 4; ++w) {
                __half h0, l0, h1, l1;
                split_half(v[u * 8 + w * 2], h0, l0);
                split_half(v[u * 8 + w * 2 + 1], h1, l1);
                hh[w] = __halves2half2(h0, h1);
                ll[w] = __halves2half2(l0, l1);
              }
              oh[u] = hv;
              ol[u] = lv;
            }
          }
          if (job.out_f32 && bg < job.B) {
            float* of = job.out_f32 + ((size_t)bg * job.Lout + lrow) * job.out_f32_C;
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const int c = n0 + cb + e;
              if (c < job.Cout) of[c] = v[e];
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc(tmem_base, job.tmem_cols);
}

cudaError_t launch_conv_job(const ConvJob& job, int n_groups, int n_slices, size_t smem_bytes, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  dim3 grid(n_groups * (kGroupSamples / job.S), n_slices);
  conv_gemm_kernel<<<grid, kConvThreads, smem_bytes, stream>>>(job);
  return cudaGetLastError();
}

cudaError_t read_device_error(unsigned int* code) {
  return cudaMemcpyFromSymbol(code, g_dm_device_error, sizeof(unsigned int));
}

}  // namespace dm
