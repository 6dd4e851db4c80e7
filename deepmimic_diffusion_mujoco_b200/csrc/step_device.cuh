// Device-side K4 (fused DDPM posterior + noise + inpaint + operand re-pack), shared by the stand-alone kernel
// (step_kernels.cu) and the persistent sampling-loop kernel (conv_kernel.cu).
// Reference: diffuser/models/diffusion.py:19-29,149-184,202-206,215; sampling_config.py:23-32.
#pragma once
#include "conv_job.h"
#include "ptx.cuh"
#include "step_kernels.h"

namespace dm {

// ------------------------------------------------------------------------------------------------
// Philox4x32-10 + Box-Muller: 4 standard normals per (counter, key)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
__device__ __forceinline__ float4 philox_normal4(unsigned long long seed, unsigned long long idx, uint32_t stream) {
  const uint4 r = philox4x32_10(make_uint4((uint32_t)idx, (uint32_t)(idx >> 32), stream, 0u),
                                make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  const float k = 2.3283064365386963e-10f;  // 2^-32
  const float u0 = ((float)r.x + 0.5f) * k, u1 = ((float)r.y + 0.5f) * k;
  const float u2 = ((float)r.z + 0.5f) * k, u3 = ((float)r.w + 0.5f) * k;
  const float ra = sqrtf(-2.f * __logf(u0)), rb = sqrtf(-2.f * __logf(u2));
  float s0, c0, s1, c1;
  __sincosf(6.283185307179586f * u1, &s0, &c0);
  __sincosf(6.283185307179586f * u3, &s1, &c1);
  return make_float4(ra * c0, ra * s0, rb * c1, rb * s1);
}

// ------------------------------------------------------------------------------------------------
// K4: one pass per denoising step over x[B,H,T].
//   x0   = predict_epsilon ? sr*x - srm1*out : out ; optional clamp(-1,1)
//   mean = c1*x0 + c2*x ; x' = mean + sigma*z*[t>0] ; x' = mask ? value : x'   (bit-exact copy)
// writes x' (fp32, public layout), the fp16 hi/lo operand planes the first conv TMA-loads, and optionally
// the chain slot.  mode 1 = init: x' = (have_start ? x : z), then inpaint (diffusion.py:202-206).
// ------------------------------------------------------------------------------------------------
// processes float4 indices [i4_begin, i4_end) with `nthreads` cooperating threads, this one being `tid`.
// coherent = true: x / model_out were written earlier in the SAME kernel by other SMs -> bypass L1.
// Conditioning operands as the kernel will read them.  With dm_cond_t::flags_dev the batch-uniform / identity / purity
// decisions are taken HERE from device-resident flags (the host passes the full [B,H*T] tensors and never reads back).
struct CondView {
  const unsigned char* mask;
  const float* value;
  long long mask_bstride, value_bstride;
  bool poison;   // the traced conditioning_fn was not a pure overwrite: the sample must not look plausible
};
__device__ __forceinline__ CondView resolve_cond(const StepParams& p) {
  CondView c{p.mask, p.value, p.mask_bstride, p.value_bstride, false};
  if (p.cond_flags != nullptr) {
    const uchar4 f = __ldg(reinterpret_cast<const uchar4*>(p.cond_flags));
    if (!f.x) c.mask = nullptr;
    if (f.y) {
      c.mask_bstride = 0;
      if (f.z) c.value_bstride = 0;
    }
    c.poison = f.w == 0;
  }
  return c;
}

template <bool kCoherent>
__device__ __forceinline__ void ddpm_step_range(const StepParams& p, long long i4_begin, long long i4_end, long long tid,
                                                long long nthreads) {
  const int HT = p.H * p.T;
  const CondView cv = resolve_cond(p);
  for (long long i4 = i4_begin + tid; i4 < i4_end; i4 += nthreads) {
    const long long e0 = i4 << 2;
    float xv[4], ov[4], zv[4];
    {
      const float4 t4 = kCoherent ? __ldcg(reinterpret_cast<const float4*>(p.x) + i4) : reinterpret_cast<const float4*>(p.x)[i4];
      xv[0] = t4.x; xv[1] = t4.y; xv[2] = t4.z; xv[3] = t4.w;
    }
    if (p.mode == 0) {
      const float4 t4 = kCoherent ? __ldcg(reinterpret_cast<const float4*>(p.model_out) + i4)
                                  : __ldg(reinterpret_cast<const float4*>(p.model_out) + i4);
      ov[0] = t4.x; ov[1] = t4.y; ov[2] = t4.z; ov[3] = t4.w;
    }
    const bool need_z = (p.mode == 0) ? (p.add_noise != 0) : (p.have_start == 0);
    if (need_z) {
      float4 t4;
      if (p.noise) t4 = __ldg(reinterpret_cast<const float4*>(p.noise) + i4);
      else t4 = philox_normal4(p.seed_ptr ? *p.seed_ptr : p.seed, p.rng_offset + (unsigned long long)i4, p.rng_stream);
      zv[0] = t4.x; zv[1] = t4.y; zv[2] = t4.z; zv[3] = t4.w;
    } else {
      zv[0] = zv[1] = zv[2] = zv[3] = 0.f;
    }
    float rv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float r;
      if (p.mode == 0) {
        float x0 = ov[k];
        if (p.predict_epsilon) x0 = __fsub_rn(__fmul_rn(p.sr, xv[k]), __fmul_rn(p.srm1, ov[k]));
        if (p.clip_denoised) x0 = fminf(fmaxf(x0, -1.f), 1.f);
        const float mean = __fadd_rn(__fmul_rn(p.c1, x0), __fmul_rn(p.c2, xv[k]));
        r = __fadd_rn(mean, __fmul_rn(p.sigma, zv[k]));
      } else {
        r = p.have_start ? xv[k] : zv[k];
      }
      rv[k] = r;
    }
    // inpaint + operand planes
    const long long b = e0 / HT;
    const int rem = (int)(e0 - b * HT);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      int r_in = rem + k;
      long long bb = b;
      if (r_in >= HT) { r_in -= HT; bb += 1; }
      if (cv.mask) {
        const size_t mi = (size_t)bb * cv.mask_bstride + r_in;
        if (cv.mask[mi]) rv[k] = cv.value[(size_t)bb * cv.value_bstride + r_in];
      }
      if (cv.poison) rv[k] = __int_as_float(0x7fc00000);
      const int h = r_in / p.T, c = r_in - h * p.T;
      const size_t row = ((size_t)(bb >> 4) * p.H + h) * kGroupSamples + (bb & 15);
      __half hi, lo;
      split_half(rv[k], hi, lo);
      p.xs[row * p.xs_C + c] = hi;
      p.xs[p.xs_plane_stride + row * p.xs_C + c] = lo;
      if (p.chain) p.chain[((size_t)bb * p.chain_len + p.chain_slot) * HT + r_in] = rv[k];
    }
    reinterpret_cast<float4*>(p.x)[i4] = make_float4(rv[0], rv[1], rv[2], rv[3]);
  }
}


}  // namespace dm
