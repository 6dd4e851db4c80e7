// The two CUDA-core stages of Residual(PreNorm(LinearAttention)) (reference: diffuser/models/helpers.py:75-131, used at
// diffuser/models/temporal.py:85,94,103).  The block runs as FOUR jobs of the layer list:
//   1. norm   (this file)   xn = (x - mean_c) / sqrt(var_c + eps) * g + b      LayerNorm over channels, per position
//   2. to_qkv (conv job)    qkv = W_qkv xn            1x1 conv, no bias, 3 x heads x 32 output channels   [tcgen05]
//   3. core   (this file)   k = softmax over positions; ctx[d][e] = sum_l k[d][l] v[e][l];
//                           o[e][l] = sum_d ctx[d][e] (q[d][l] * 32^-0.5)
//   4. to_out (conv job)    y = W_out o + b_out + x   1x1 conv with bias and the block's residual          [tcgen05]
// so the two 1x1 convs (two thirds of the block's FLOPs) run on the tensor pipe through the ordinary conv path, and all
// four run inside the persistent sampling loop (conv_kernel.cu) as well as one launch each (attn_kernel.cu).
// Both stages are called by the 16 epilogue warps (512 threads) of a CTA for its 8 samples of one 16-sample group and go
// global -> global (activation planes [2][G][L][16][C], value = hi + lo; the working set is L2 resident).
// The descriptor is a ConvJob with `kind` != 0; the fields it uses:
//   res / res_C / res_plane_stride   input planes      out / out_C / out_plane_stride   output planes
//   Lj   positions                    gamma / beta      LayerNorm g / b (norm stage)
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

#include "conv_job.h"

namespace dm {

constexpr int kAttnHeads = 4;        // LinearAttention defaults (helpers.py:111); temporal.py never overrides them
constexpr int kAttnDh = 32;
constexpr int kAttnHid = kAttnHeads * kAttnDh;
constexpr int kCoreRows = 16;        // positions staged per chunk by the core stage
constexpr int kCoreWarpFloats = 4 * kCoreRows * 32;   // per warp: k and v chunks of 32 channels, double-buffered [2][16][32] each
constexpr size_t kCoreSmemBytes = (size_t)16 * kCoreWarpFloats * sizeof(float);

// L2-coherent loads throughout: inside the persistent kernel these bytes were written moments ago (by this SM's TMA stores
// or generic stores), and a stale L1 line must not be read.
__device__ __forceinline__ float ld_hl(const __half* hi, long long plane) {
  const unsigned short h = __ldcg(reinterpret_cast<const unsigned short*>(hi));
  const unsigned short l = __ldcg(reinterpret_cast<const unsigned short*>(hi + plane));
  return __half2float(__ushort_as_half(h)) + __half2float(__ushort_as_half(l));
}
__device__ __forceinline__ void st_hl(__half* hi, long long plane, float v) {
  const __half h = __float2half_rn(v);
  hi[0] = h;
  hi[plane] = __float2half_rn(v - __half2float(h));
}
// 8 consecutive channels (16 bytes per plane): value = hi + lo
__device__ __forceinline__ void ld_hl8(const __half* hi, long long plane, float* x) {
  const uint4 h = __ldcg(reinterpret_cast<const uint4*>(hi));
  const uint4 l = __ldcg(reinterpret_cast<const uint4*>(hi + plane));
  const __half2* hh = reinterpret_cast<const __half2*>(&h);
  const __half2* ll = reinterpret_cast<const __half2*>(&l);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 a = __half22float2(hh[i]), b = __half22float2(ll[i]);
    x[2 * i] = a.x + b.x;
    x[2 * i + 1] = a.y + b.y;
  }
}
__device__ __forceinline__ void st_hl8(__half* hi, long long plane, const float* x) {
  uint4 h, l;
  __half2* hh = reinterpret_cast<__half2*>(&h);
  __half2* ll = reinterpret_cast<__half2*>(&l);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    hh[i] = __floats2half2_rn(x[2 * i], x[2 * i + 1]);
    const float2 b = __half22float2(hh[i]);
    ll[i] = __floats2half2_rn(x[2 * i] - b.x, x[2 * i + 1] - b.y);
  }
  *reinterpret_cast<uint4*>(hi) = h;
  *reinterpret_cast<uint4*>(hi + plane) = l;
}

// ---- stage 1: LayerNorm over channels (helpers.py:84-95: biased variance, eps inside the sqrt) ----
// C / 8 lanes (at most 32) share a row, 8 channels (16 bytes per plane) each per 256-channel segment; a warp works on
// 32 / lanes rows at a time and a row makes ONE trip through L2 (the values stay in registers between the passes).
__device__ __forceinline__ void attn_norm_stage(const ConvJob& job, int g, int b0, int tid) {
  const int C = job.res_C, L = job.Lj;
  const int warp = tid >> 5, lane = tid & 31;
  const float inv_c = 1.f / (float)C;
  const long long ip = job.res_plane_stride, op = job.out_plane_stride;
  if ((C & (C - 1)) == 0 && C >= 32 && C <= 1024) {
    const int lpr = C >= 256 ? 32 : C >> 3;      // lanes per row
    const int rpw = 32 / lpr;                    // rows per warp and iteration
    const int nseg = (C + 255) >> 8;             // 256-channel segments (1..4)
    const int sub = lane / lpr, li = lane - sub * lpr;
    for (int r = warp * rpw + sub; r < L * 8; r += 16 * rpw) {
      const int l = r >> 3, b = r & 7;
      const size_t row = ((size_t)g * L + l) * kGroupSamples + b0 + b;
      const __half* __restrict__ in = job.res + row * C + li * 8;
      __half* __restrict__ out = job.out + row * job.out_C + li * 8;
      float x[4][8];
      float s = 0.f;
#pragma unroll
      for (int sg = 0; sg < 4; ++sg)
        if (sg < nseg) {
          ld_hl8(in + sg * 256, ip, x[sg]);
#pragma unroll
          for (int i = 0; i < 8; ++i) s += x[sg][i];
        }
      for (int o = lpr >> 1; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const float mean = s * inv_c;
      float q = 0.f;
#pragma unroll
      for (int sg = 0; sg < 4; ++sg)
        if (sg < nseg) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { x[sg][i] -= mean; q = fmaf(x[sg][i], x[sg][i], q); }
        }
      for (int o = lpr >> 1; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
      const float inv = 1.f / sqrtf(q * inv_c + 1e-5f);
#pragma unroll
      for (int sg = 0; sg < 4; ++sg)
        if (sg < nseg) {
          const float4* gp = reinterpret_cast<const float4*>(job.gamma + sg * 256 + li * 8);
          const float4* bp = reinterpret_cast<const float4*>(job.beta + sg * 256 + li * 8);
          const float4 g0 = __ldg(gp), g1 = __ldg(gp + 1), e0 = __ldg(bp), e1 = __ldg(bp + 1);
          float y[8];
          y[0] = x[sg][0] * inv * g0.x + e0.x; y[1] = x[sg][1] * inv * g0.y + e0.y;
          y[2] = x[sg][2] * inv * g0.z + e0.z; y[3] = x[sg][3] * inv * g0.w + e0.w;
          y[4] = x[sg][4] * inv * g1.x + e1.x; y[5] = x[sg][5] * inv * g1.y + e1.y;
          y[6] = x[sg][6] * inv * g1.z + e1.z; y[7] = x[sg][7] * inv * g1.w + e1.w;
          st_hl8(out + sg * 256, op, y);
        }
    }
    return;
  }
  // any other channel count: one warp per row, three trips through L2
  for (int r = warp; r < L * 8; r += 16) {
    const int l = r >> 3, b = r & 7;
    const size_t row = ((size_t)g * L + l) * kGroupSamples + b0 + b;
    const __half* in = job.res + row * C;
    __half* out = job.out + row * job.out_C;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += ld_hl(in + c, ip);
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * inv_c;
    float q = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float d = ld_hl(in + c, ip) - mean;
      q = fmaf(d, d, q);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float inv = 1.f / sqrtf(q * inv_c + 1e-5f);
    for (int c = lane; c < C; c += 32)
      st_hl(out + c, op, (ld_hl(in + c, ip) - mean) * inv * __ldg(job.gamma + c) + __ldg(job.beta + c));
  }
}

// ---- stage 3: softmax over positions, context, output.  One warp per (sample, head). ----
// Input: the to_qkv layer's activation planes (384 channels: q | k | v = [0,128) | [128,256) | [256,384)).
// The positions go through shared memory kCoreRows at a time with cp.async (raw hi / lo halves, 16 bytes per lane, no
// registers), double-buffered: chunk c + 1 is in flight while chunk c is converted to fp32 in place and consumed.  The softmax is accumulated on line
// (running maximum; context and denominator rescaled when it grows), so k, v and q are read once.
// sm: kCoreSmemBytes of shared memory, 16-byte aligned.
__device__ __forceinline__ void cp_async16(float* dst_smem, const float* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// n (<= kCoreRows) rows x 32 channels of one head, raw: tile row = 64 bytes of the hi plane | 64 bytes of the lo plane.
// 8 lanes per row (4 x 16 bytes per plane), 4 rows per instruction.
__device__ __forceinline__ void core_issue_tile(float* tile, const __half* base, long long plane, size_t rstep, int n, int lane) {
  const int sub = lane >> 3, pl = (lane >> 2) & 1, piece = lane & 3;
  const __half* src = base + (pl ? plane : 0) + piece * 8;
  float* dst = tile + pl * 16 + piece * 4;
#pragma unroll
  for (int k = 0; k < kCoreRows / 4; ++k) {
    const int i = k * 4 + sub;
    if (i < n) cp_async16(dst + i * 32, reinterpret_cast<const float*>(src + (size_t)i * rstep));
  }
}
// In place: raw rows (hi | lo halves) -> 32 floats (hi + lo).  8 lanes read one row's 128 bytes (16 each: no bank
// conflicts), the lo pieces move to the lanes that hold the matching hi pieces by shuffle, and those lanes write 8 floats
// each.  Every lane reads before any lane writes.
__device__ __forceinline__ void core_convert_tile(float* tile, int n, int lane) {
  const int sub = lane >> 3, j = lane & 7;
  uint4 raw[kCoreRows / 4];
#pragma unroll
  for (int k = 0; k < kCoreRows / 4; ++k) {
    const int row = k * 4 + sub;
    raw[k] = row < n ? *reinterpret_cast<const uint4*>(tile + row * 32 + j * 4) : make_uint4(0u, 0u, 0u, 0u);
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < kCoreRows / 4; ++k) {
    const int row = k * 4 + sub;
    uint4 lo;
    lo.x = __shfl_down_sync(0xffffffffu, raw[k].x, 4);
    lo.y = __shfl_down_sync(0xffffffffu, raw[k].y, 4);
    lo.z = __shfl_down_sync(0xffffffffu, raw[k].z, 4);
    lo.w = __shfl_down_sync(0xffffffffu, raw[k].w, 4);
    if (j < 4 && row < n) {
      const __half2* hh = reinterpret_cast<const __half2*>(&raw[k]);
      const __half2* ll = reinterpret_cast<const __half2*>(&lo);
      float x[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 a = __half22float2(hh[i]), b = __half22float2(ll[i]);
        x[2 * i] = a.x + b.x;
        x[2 * i + 1] = a.y + b.y;
      }
      float4* d = reinterpret_cast<float4*>(tile + row * 32 + j * 8);
      d[0] = make_float4(x[0], x[1], x[2], x[3]);
      d[1] = make_float4(x[4], x[5], x[6], x[7]);
    }
  }
  __syncwarp();
}

__device__ __forceinline__ void attn_core_stage(const ConvJob& job, int g, int b0, int tid, float* sm) {
  const int L = job.Lj, C3 = job.res_C, CO = job.out_C;
  const long long ip = job.res_plane_stride, op = job.out_plane_stride;
  const int warp = tid >> 5, lane = tid & 31;
  float* kb = sm + warp * kCoreWarpFloats;       // [2][kCoreRows][32]: k chunks (softmax numerators in place), later q chunks
  float* vb = kb + 2 * kCoreRows * 32;           // [2][kCoreRows][32]: v chunks
  const float scale = 0.17677669529663687f;      // 32^-0.5
  const size_t rstep = (size_t)kGroupSamples * C3;   // halfs between positions of one sample
  const int nchunk = (L + kCoreRows - 1) / kCoreRows;
  const int d0 = (lane >> 2) * 4, e0 = (lane & 3) * 8;
  for (int pair = warp; pair < 8 * kAttnHeads; pair += 16) {
    const int b = pair >> 2, h = pair & 3;
    const size_t row0 = (size_t)g * L * kGroupSamples + b0 + b;           // row of position l: row0 + l * 16
    const __half* qp = job.res + row0 * C3 + h * kAttnDh;                 // q | k | v = channels [0,128) | [128,256) | [256,384)
    const __half* kp = qp + kAttnHid;
    const __half* vp = kp + kAttnHid;
    // Register blocking: lane (dg, eg) owns ctx[4 dg .. 4 dg + 3][8 eg .. 8 eg + 7].  A row then costs 3 x 16 bytes of shared
    // memory per lane (4 numerators, 8 v values) for 32 FMAs; one lane per d would need all 32 v values (the shared-memory
    // pipe delivers 128 bytes per cycle to the whole SM, broadcast or not: that was the bound, 4x below the FMA rate).
    float acc[4][8];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[a][e] = 0.f;
    float m = -3.0e38f, sum = 0.f;                                        // softmax statistics of k[d][.]: lane = d
    core_issue_tile(kb, kp, ip, rstep, min(kCoreRows, L), lane);
    core_issue_tile(vb, vp, ip, rstep, min(kCoreRows, L), lane);
    cp_async_commit();
    for (int c = 0; c < nchunk; ++c) {
      const int n = min(kCoreRows, L - c * kCoreRows);
      if (c + 1 < nchunk) {
        const int l1 = (c + 1) * kCoreRows, n1 = min(kCoreRows, L - l1);
        core_issue_tile(kb + ((c + 1) & 1) * kCoreRows * 32, kp + (size_t)l1 * rstep, ip, rstep, n1, lane);
        core_issue_tile(vb + ((c + 1) & 1) * kCoreRows * 32, vp + (size_t)l1 * rstep, ip, rstep, n1, lane);
      } else {
        core_issue_tile(kb + ((c + 1) & 1) * kCoreRows * 32, qp, ip, rstep, min(kCoreRows, L), lane);   // first q chunk
      }
      cp_async_commit();
      cp_async_wait<1>();
      __syncwarp();
      float* kt = kb + (c & 1) * kCoreRows * 32;
      float* vt = vb + (c & 1) * kCoreRows * 32;
      core_convert_tile(kt, n, lane);
      core_convert_tile(vt, n, lane);
      float mc = m;
      for (int i = 0; i < n; ++i) mc = fmaxf(mc, kt[i * 32 + lane]);
      const float rs = expf(m - mc);
      m = mc;
      sum *= rs;
      // the numerators, written over k in place (a lane only touches its own column here)
#pragma unroll 4
      for (int i = 0; i < n; ++i) {
        const float p = expf(kt[i * 32 + lane] - m);
        kt[i * 32 + lane] = p;
        sum += p;
      }
      __syncwarp();
#pragma unroll
      for (int a = 0; a < 4; ++a) {                                       // the running maximum of row d grew: rescale
        const float ra = __shfl_sync(0xffffffffu, rs, d0 + a);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[a][e] *= ra;
      }
#pragma unroll 2
      for (int i = 0; i < n; ++i) {
        const float4 p4 = *reinterpret_cast<const float4*>(kt + i * 32 + d0);
        const float4 va = *reinterpret_cast<const float4*>(vt + i * 32 + e0);
        const float4 vb4 = *reinterpret_cast<const float4*>(vt + i * 32 + e0 + 4);
        const float pv[4] = {p4.x, p4.y, p4.z, p4.w};
        const float vv[8] = {va.x, va.y, va.z, va.w, vb4.x, vb4.y, vb4.z, vb4.w};
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[a][e] = fmaf(pv[a], vv[e], acc[a][e]);
      }
      __syncwarp();
    }
    // ctx[d][e] / sum_d, with q's scale folded in
    {
      const float f = scale / sum;
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const float fa = __shfl_sync(0xffffffffu, f, d0 + a);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[a][e] *= fa;
      }
    }
    // o[e][l] = sum_d ctx[d][e] q[d][l]: every lane adds its 4 d's for its 8 e's; the 8 lanes that share an e group then
    // reduce-scatter (xor 16, 8, 4), and each lane is left with ONE finished output: e = e0 + 4 * bit4 + 2 * bit3 + bit2.
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    const int e_out = e0 + (b4 ? 4 : 0) + (b3 ? 2 : 0) + (b2 ? 1 : 0);
    __half* outp = job.out + ((size_t)g * L * kGroupSamples + b0 + b) * CO + h * kAttnDh + e_out;   // 'b h c d -> b (h c) d'
    const size_t ostep = (size_t)kGroupSamples * CO;
    // q chunk c sits in kb[(nchunk + c) & 1] (chunk 0 was issued with the last k / v chunk)
    for (int c = 0; c < nchunk; ++c) {
      const int n = min(kCoreRows, L - c * kCoreRows);
      if (c + 1 < nchunk) {
        const int l1 = (c + 1) * kCoreRows;
        core_issue_tile(kb + ((nchunk + c + 1) & 1) * kCoreRows * 32, qp + (size_t)l1 * rstep, ip, rstep, min(kCoreRows, L - l1), lane);
      }
      cp_async_commit();
      cp_async_wait<1>();
      __syncwarp();
      float* qt = kb + ((nchunk + c) & 1) * kCoreRows * 32;
      core_convert_tile(qt, n, lane);
#pragma unroll 2
      for (int i = 0; i < n; ++i) {
        const float4 q4 = *reinterpret_cast<const float4*>(qt + i * 32 + d0);
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = fmaf(acc[3][e], q4.w, fmaf(acc[2][e], q4.z, fmaf(acc[1][e], q4.y, acc[0][e] * q4.x)));
        float r4[4], r2[2];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float recv = __shfl_xor_sync(0xffffffffu, b4 ? o[t] : o[t + 4], 16);
          r4[t] = (b4 ? o[t + 4] : o[t]) + recv;
        }
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          const float recv = __shfl_xor_sync(0xffffffffu, b3 ? r4[t] : r4[t + 2], 8);
          r2[t] = (b3 ? r4[t + 2] : r4[t]) + recv;
        }
        const float recv = __shfl_xor_sync(0xffffffffu, b2 ? r2[0] : r2[1], 4);
        st_hl(outp + (size_t)(c * kCoreRows + i) * ostep, op, (b2 ? r2[1] : r2[0]) + recv);
      }
      __syncwarp();
    }
    cp_async_wait<0>();
  }
}

}  // namespace dm
