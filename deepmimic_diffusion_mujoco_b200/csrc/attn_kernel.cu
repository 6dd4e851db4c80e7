// Residual(PreNorm(LinearAttention)) of the TemporalUnet's attention=True variant
// (reference: diffuser/models/helpers.py:75-131, used at diffuser/models/temporal.py:85,94,103):
//   xn  = (x - mean_c) / sqrt(var_c + eps) * g + b                       LayerNorm over channels, per position
//   qkv = W_qkv xn            (1x1 conv, no bias; 3 x heads x 32 rows)
//   k   = softmax over positions;  q = q * 32^-0.5
//   ctx[d][e] = sum_l k[d][l] v[e][l];   o[e][l] = sum_d ctx[d][e] q[d][l]
//   y   = W_out o + b_out + x
// One CTA owns NS samples of one 16-sample activation group and keeps everything in shared memory: the
// activation tile is read once (fp16 hi + lo planes, value = hi + lo) and written once.  All arithmetic is fp32
// on the CUDA cores: the contractions are 32-wide per head and the whole op is ~17 % of the U-Net's FLOPs, so this
// first version favours exactness over tensor-core use; it runs in the per-layer launch path only (the fused
// persistent loop kernel is used when attention=False).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "attn_kernel.h"
#include "conv_job.h"

namespace dm {

constexpr int kAttnThreads = 512;
constexpr int kLc = 8;   // positions per register tile

__device__ __forceinline__ int ceil_div(int a, int b) { return (a + b - 1) / b; }

__host__ __device__ __forceinline__ int row_pitch(int L) { return (L & 3) ? (L | 1) : L + 4; }

// kVec: L is a multiple of 4 -> rows are 16-byte aligned (pitch L + 4) and the two 1x1 convs run on 4 x 4 register
// tiles fed by one 16-byte weight load and one 16-byte activation load per 16 FMAs.
template <bool kVec>
__global__ void __launch_bounds__(kAttnThreads, 2) linear_attention_kernel(const AttnParams p) {
  extern __shared__ __align__(16) float sm[];
  const int C = p.C, L = p.L, NS = p.NS;
  const int Lp = row_pitch(L);
  float* xs = sm;                             // [NS][C][Lp]  input, becomes the output accumulator
  float* xn = xs + NS * C * Lp;               // [NS][C][Lp]  normalised input
  float* qkv = xn + NS * C * Lp;              // [NS][96][Lp] q | k (later o) | v of the current head
  float* ctx = qkv + NS * 96 * Lp;            // [NS][32][33]
  float* wsm = ctx + NS * 32 * 33 + ((NS * 32 * 33) & 3 ? 4 - ((NS * 32 * 33) & 3) : 0);   // [2][32][96] weight chunks (16-byte aligned, double-buffered)
  const int tid = threadIdx.x;
  const int per_group = kGroupSamples / NS;
  const int g = blockIdx.x / per_group;
  const int s0 = (blockIdx.x % per_group) * NS;

  // ---- load: activation layout [plane][G][L][16][C] ----
  for (int i = tid; i < NS * L * C; i += kAttnThreads) {
    const int c = i % C, l = (i / C) % L, s = i / (C * L);
    const long long off = (((long long)g * L + l) * kGroupSamples + s0 + s) * C + c;
    xs[(s * C + c) * Lp + l] = __half2float(p.in[off]) + __half2float(p.in[p.in_plane + off]);
  }
  __syncthreads();
  // ---- LayerNorm over channels (helpers.py:84-95: biased variance, eps inside the sqrt) ----
  for (int i = tid; i < NS * L; i += kAttnThreads) {
    const int l = i % L, s = i / L;
    const float* col = xs + s * C * Lp + l;
    float mean = 0.f;
    for (int c = 0; c < C; ++c) mean += col[c * Lp];
    mean /= (float)C;
    float var = 0.f;
    for (int c = 0; c < C; ++c) { const float d = col[c * Lp] - mean; var = fmaf(d, d, var); }
    var /= (float)C;
    const float inv = 1.f / sqrtf(var + p.eps);
    float* dst = xn + s * C * Lp + l;
    for (int c = 0; c < C; ++c) dst[c * Lp] = (col[c * Lp] - mean) * inv * __ldg(p.ln_g + c) + __ldg(p.ln_b + c);
  }
  __syncthreads();
  const int n_lc = ceil_div(L, kLc);
  for (int h = 0; h < p.heads; ++h) {
    // ---- q, k, v rows of this head: W_qkv^T is [C][3*heads*32] so that a warp reads consecutive rows coalesced ----
    if (kVec) {
      // The head's weight rows are staged through shared memory 32 input channels at a time by ALL threads (one
      // exposed L2 latency per chunk instead of one per channel: at C = 256, L = 8 only 192 threads own a tile).
      const int n_l4 = L >> 2;
      const int OW = 3 * p.heads * 32;
      const int n_items = 24 * NS * n_l4;
      for (int i0 = 0; i0 < n_items; i0 += kAttnThreads) {
        const int i = i0 + tid;
        const bool live = i < n_items;
        const int oq = i % 24, s = live ? (i / 24) % NS : 0, l0 = live ? (i / (24 * NS)) << 2 : 0;
        const int o = oq << 2;
        const float* xr = xn + s * C * Lp + l0;
        float acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
        // double-buffered cp.async staging: chunk cb + 32 is in flight while chunk cb is consumed
        auto stage = [&](int cb, float* dst) {
#pragma unroll
          for (int k = 0; k < (32 * 96) / kAttnThreads; ++k) {
            const int j = tid + k * kAttnThreads;
            const int cc = j / 96, r = j - cc * 96;
            const float* src = p.wqkv_t + (size_t)(cb + cc) * OW + (r >> 5) * (p.heads * 32) + h * 32 + (r & 31);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + j)), "l"(src) : "memory");
          }
          asm volatile("cp.async.commit_group;" ::: "memory");
        };
        __syncthreads();            // both buffers free (previous round / head fully consumed)
        stage(0, wsm);
        int buf = 0;
        for (int cb = 0; cb < C; cb += 32, buf ^= 1) {
          const float* wcur = wsm + buf * (32 * 96);
          if (cb + 32 < C) {
            stage(cb + 32, wsm + (buf ^ 1) * (32 * 96));
            asm volatile("cp.async.wait_group 1;" ::: "memory");
          } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
          }
          __syncthreads();
          if (live) {
#pragma unroll 4
            for (int cc = 0; cc < 32; ++cc) {
              const float4 w = *reinterpret_cast<const float4*>(wcur + cc * 96 + o);
              const float4 x = *reinterpret_cast<const float4*>(xr + (cb + cc) * Lp);
              const float wv[4] = {w.x, w.y, w.z, w.w}, xv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
              for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(wv[a], xv[b], acc[a][b]);
            }
          }
          __syncthreads();          // chunk consumed before its buffer is refilled two iterations later
        }
        if (live) {
          const float sc = (o >> 5) == 0 ? p.scale : 1.f;
#pragma unroll
          for (int a = 0; a < 4; ++a)
            *reinterpret_cast<float4*>(qkv + (s * 96 + o + a) * Lp + l0) =
                make_float4(acc[a][0] * sc, acc[a][1] * sc, acc[a][2] * sc, acc[a][3] * sc);
        }
      }
    } else
    for (int i = tid; i < 96 * NS * n_lc; i += kAttnThreads) {
      const int o = i % 96, s = (i / 96) % NS, lc = i / (96 * NS);
      const int part = o >> 5, d = o & 31;
      const int row = part * (p.heads * 32) + h * 32 + d;
      const int l0 = lc * kLc;
      float acc[kLc];
#pragma unroll
      for (int j = 0; j < kLc; ++j) acc[j] = 0.f;
      const float* xr = xn + s * C * Lp + l0;
      const float* wr = p.wqkv_t + row;
      const int OW = 3 * p.heads * 32;
      if (l0 + kLc <= L) {
        for (int c = 0; c < C; ++c) {
          const float w = __ldg(wr + (size_t)c * OW);
#pragma unroll
          for (int j = 0; j < kLc; ++j) acc[j] = fmaf(w, xr[c * Lp + j], acc[j]);
        }
      } else {
        for (int c = 0; c < C; ++c) {
          const float w = __ldg(wr + (size_t)c * OW);
#pragma unroll
          for (int j = 0; j < kLc; ++j) if (l0 + j < L) acc[j] = fmaf(w, xr[c * Lp + j], acc[j]);
        }
      }
      const float sc = part == 0 ? p.scale : 1.f;
      float* dst = qkv + (s * 96 + o) * Lp + l0;
#pragma unroll
      for (int j = 0; j < kLc; ++j) if (l0 + j < L) dst[j] = acc[j] * sc;
    }
    __syncthreads();
    // ---- softmax of every k row over the positions ----
    for (int i = tid; i < NS * 32; i += kAttnThreads) {
      float* row = qkv + ((i >> 5) * 96 + 32 + (i & 31)) * Lp;
      float m = row[0];
      for (int l = 1; l < L; ++l) m = fmaxf(m, row[l]);
      float sum = 0.f;
      for (int l = 0; l < L; ++l) { const float e = expf(row[l] - m); row[l] = e; sum += e; }
      const float inv = 1.f / sum;
      for (int l = 0; l < L; ++l) row[l] *= inv;
    }
    __syncthreads();
    // ---- ctx[d][e] = sum_l k[d][l] v[e][l] ----
    if (kVec) {
      // 4 x 4 register tile over rows {dq + 8a} x {eq + 8b}: rows 8 apart keep the 16-byte row loads of the eight
      // threads that share a wavefront on distinct banks (pitch = 4 mod 32 words)
      for (int i = tid; i < NS * 64; i += kAttnThreads) {
        const int eq = i & 7, dq = (i >> 3) & 7, s = i >> 6;
        const float* kr = qkv + (s * 96 + 32 + dq) * Lp;
        const float* vr = qkv + (s * 96 + 64 + eq) * Lp;
        float acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
        for (int l = 0; l < L; l += 4) {
          float4 kv[4], vv[4];
#pragma unroll
          for (int a = 0; a < 4; ++a) kv[a] = *reinterpret_cast<const float4*>(kr + a * 8 * Lp + l);
#pragma unroll
          for (int b = 0; b < 4; ++b) vv[b] = *reinterpret_cast<const float4*>(vr + b * 8 * Lp + l);
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b)
              acc[a][b] = fmaf(kv[a].x, vv[b].x, fmaf(kv[a].y, vv[b].y, fmaf(kv[a].z, vv[b].z, fmaf(kv[a].w, vv[b].w, acc[a][b]))));
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) ctx[(s * 32 + dq + 8 * a) * 33 + eq + 8 * b] = acc[a][b];
      }
    } else
    for (int i = tid; i < NS * 1024; i += kAttnThreads) {
      const int e = i & 31, d = (i >> 5) & 31, s = i >> 10;
      const float* kr = qkv + (s * 96 + 32 + d) * Lp;
      const float* vr = qkv + (s * 96 + 64 + e) * Lp;
      float a = 0.f;
      for (int l = 0; l < L; ++l) a = fmaf(kr[l], vr[l], a);
      ctx[(s * 32 + d) * 33 + e] = a;
    }
    __syncthreads();
    // ---- o[e][l] = sum_d ctx[d][e] q[d][l]; stored over the (dead) k rows ----
    if (kVec) {
      const int n_l4 = L >> 2;
      for (int i = tid; i < NS * 8 * n_l4; i += kAttnThreads) {
        const int eq = i & 7, l0 = ((i >> 3) % n_l4) << 2, s = (i >> 3) / n_l4;
        const float* qc = qkv + (s * 96) * Lp + l0;
        const float* cc = ctx + (s * 32) * 33 + eq;
        float acc[4][4];
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[b][j] = 0.f;
#pragma unroll 4
        for (int d = 0; d < 32; ++d) {
          const float4 q4 = *reinterpret_cast<const float4*>(qc + d * Lp);
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const float cv = cc[d * 33 + 8 * b];
            acc[b][0] = fmaf(cv, q4.x, acc[b][0]); acc[b][1] = fmaf(cv, q4.y, acc[b][1]);
            acc[b][2] = fmaf(cv, q4.z, acc[b][2]); acc[b][3] = fmaf(cv, q4.w, acc[b][3]);
          }
        }
#pragma unroll
        for (int b = 0; b < 4; ++b)   // k rows are dead after the ctx pass (barrier above)
          *reinterpret_cast<float4*>(qkv + (s * 96 + 32 + eq + 8 * b) * Lp + l0) = make_float4(acc[b][0], acc[b][1], acc[b][2], acc[b][3]);
      }
    } else
    for (int i = tid; i < NS * 32 * L; i += kAttnThreads) {
      const int l = i % L, e = (i / L) & 31, s = i / (32 * L);
      const float* qc = qkv + (s * 96) * Lp + l;
      const float* cc = ctx + (s * 32) * 33 + e;
      float a = 0.f;
#pragma unroll 8
      for (int d = 0; d < 32; ++d) a = fmaf(cc[d * 33], qc[d * Lp], a);
      qkv[(s * 96 + 32 + e) * Lp + l] = a;   // k rows of this (s, e, l) are only read in the ctx pass above
    }
    __syncthreads();
    // ---- xs += W_out[:, head] o  (+ bias once): W_out^T is [heads*32][C] ----
    if (kVec) {
      // W_out^T rows of this head ([32][C]) are staged through the same double buffer, ec rows (ec * C <= 3072) at a time
      const int n_l4 = L >> 2, n_c4 = C >> 2;
      const int ec = C <= 96 ? 32 : (C <= 192 ? 16 : (C <= 384 ? 8 : (C <= 768 ? 4 : (C <= 1536 ? 2 : 1))));
      const int n_items = n_c4 * NS * n_l4;
      auto stage_o = [&](int e0, float* dst) {
        for (int j = tid; j < ec * C; j += kAttnThreads) {
          const float* src = p.wout_t + (size_t)(h * 32 + e0) * C + j;
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + j)), "l"(src) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
      };
      for (int i0 = 0; i0 < n_items; i0 += kAttnThreads) {
        const int i = i0 + tid;
        const bool live = i < n_items;
        const int co = (i % n_c4) << 2, s = live ? (i / n_c4) % NS : 0, l0 = live ? (i / (n_c4 * NS)) << 2 : 0;
        float acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          const float b0 = (h == 0 && live) ? __ldg(p.bout + co + a) : 0.f;
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = b0;
        }
        const float* orow = qkv + (s * 96 + 32) * Lp + l0;
        __syncthreads();            // both buffers free
        stage_o(0, wsm);
        int buf = 0;
        for (int e0 = 0; e0 < 32; e0 += ec, buf ^= 1) {
          const float* wcur = wsm + buf * (32 * 96);
          if (e0 + ec < 32) {
            stage_o(e0 + ec, wsm + (buf ^ 1) * (32 * 96));
            asm volatile("cp.async.wait_group 1;" ::: "memory");
          } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
          }
          __syncthreads();
          if (live) {
#pragma unroll 4
            for (int e = 0; e < ec; ++e) {
              const float4 w = *reinterpret_cast<const float4*>(wcur + e * C + co);
              const float4 x = *reinterpret_cast<const float4*>(orow + (e0 + e) * Lp);
              const float wv[4] = {w.x, w.y, w.z, w.w}, xv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
              for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(wv[a], xv[b], acc[a][b]);
            }
          }
          __syncthreads();
        }
        if (live) {
#pragma unroll
          for (int a = 0; a < 4; ++a) {
            float4* dst = reinterpret_cast<float4*>(xs + (s * C + co + a) * Lp + l0);
            float4 v = *dst;
            v.x += acc[a][0]; v.y += acc[a][1]; v.z += acc[a][2]; v.w += acc[a][3];
            *dst = v;
          }
        }
      }
    } else
    for (int i = tid; i < C * NS * n_lc; i += kAttnThreads) {
      const int co = i % C, s = (i / C) % NS, lc = i / (C * NS);
      const int l0 = lc * kLc;
      float acc[kLc];
      const float b0 = h == 0 ? __ldg(p.bout + co) : 0.f;
#pragma unroll
      for (int j = 0; j < kLc; ++j) acc[j] = b0;
      const float* orow = qkv + (s * 96 + 32) * Lp + l0;
      const float* wr = p.wout_t + (size_t)(h * 32) * C + co;
      for (int e = 0; e < 32; ++e) {
        const float w = __ldg(wr + (size_t)e * C);
#pragma unroll
        for (int j = 0; j < kLc; ++j) if (l0 + j < L) acc[j] = fmaf(w, orow[e * Lp + j], acc[j]);
      }
      float* dst = xs + (s * C + co) * Lp + l0;
#pragma unroll
      for (int j = 0; j < kLc; ++j) if (l0 + j < L) dst[j] += acc[j];
    }
    __syncthreads();
  }
  // ---- store: split to fp16 hi + lo planes ----
  for (int i = tid; i < NS * L * C; i += kAttnThreads) {
    const int c = i % C, l = (i / C) % L, s = i / (C * L);
    const long long off = (((long long)g * L + l) * kGroupSamples + s0 + s) * C + c;
    const float v = xs[(s * C + c) * Lp + l];
    const __half hi = __float2half_rn(v);
    p.out[off] = hi;
    p.out[p.out_plane + off] = __float2half_rn(v - __half2float(hi));
  }
}

size_t attention_smem_bytes(int C, int L, int NS) {
  const int Lp = row_pitch(L);
  return sizeof(float) * ((size_t)2 * NS * C * Lp + (size_t)NS * 96 * Lp + (size_t)NS * 32 * 33 + 4 + 2 * 32 * 96);
}

int attention_pick_ns(int C, int L) {
  for (int ns = 4; ns >= 1; ns >>= 1)
    if (attention_smem_bytes(C, L, ns) <= 224 * 1024) return ns;
  return 0;
}

cudaError_t launch_linear_attention(const AttnParams& p, int n_groups, cudaStream_t s) {
  const size_t smem = attention_smem_bytes(p.C, p.L, p.NS);
  auto kern = (p.L & 3) ? linear_attention_kernel<false> : linear_attention_kernel<true>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  kern<<<n_groups * (kGroupSamples / p.NS), kAttnThreads, smem, s>>>(p);
  return cudaGetLastError();
}

}  // namespace dm
