// K3 (time-embedding rows), K4 (fused DDPM posterior + noise + inpaint + operand re-pack) and the layout
// pack/unpack helpers.  All HBM-bound elementwise work: coalesced, float4-vectorised where alignment allows.
//
// Reference restated:
//   SinusoidalPosEmb / time_mlp      diffuser/models/helpers.py:17-29, temporal.py:67-72, :28-32
//   default_sample_fn                diffuser/models/diffusion.py:19-29
//   predict_start_from_noise         diffuser/models/diffusion.py:149-160
//   q_posterior / p_mean_variance    diffuser/models/diffusion.py:162-184
//   apply_conditioning (inpaint)     diffuser/models/sampling_config.py:23-32 and the notebook closures
#include "conv_job.h"
#include "ptx.cuh"
#include "step_kernels.h"
#include "step_device.cuh"

namespace dm {

// ------------------------------------------------------------------------------------------------
// K3: time embedding -> per-block time-bias rows.  One CTA per distinct timestep value.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float mish_accurate(float x) {
  // torch.nn.Mish: x * tanh(softplus(x)), softplus threshold 20
  float sp = x > 20.f ? x : log1pf(expf(x));
  return x * tanhf(sp);
}

__global__ void time_bias_kernel(TimeMlpParams p, const float* __restrict__ t, float* __restrict__ out) {
  extern __shared__ float sm[];
  float* emb = sm;                 // [D]
  float* h1 = emb + p.D;           // [4D]
  float* m = h1 + 4 * p.D;         // [D]
  const int row = blockIdx.x;
  const float tv = t[row];
  const int half = p.D / 2;
  const float scale = -(float)(9.210340371976184 / (double)(half - 1));  // -log(10000)/(half-1)
  for (int i = threadIdx.x; i < half; i += blockDim.x) {
    const float f = expf((float)i * scale);
    const float a = tv * f;
    emb[i] = sinf(a);
    emb[half + i] = cosf(a);
  }
  __syncthreads();
  for (int o = threadIdx.x; o < 4 * p.D; o += blockDim.x) {
    float acc = p.b1[o];
    for (int k = 0; k < p.D; ++k) acc = fmaf(p.w1[o * p.D + k], emb[k], acc);
    h1[o] = mish_accurate(acc);
  }
  __syncthreads();
  for (int o = threadIdx.x; o < p.D; o += blockDim.x) {
    float acc = p.b2[o];
    for (int k = 0; k < 4 * p.D; ++k) acc = fmaf(p.w2[o * 4 * p.D + k], h1[k], acc);
    m[o] = mish_accurate(acc);   // every block's time_mlp starts with Mish (temporal.py:28-29)
  }
  __syncthreads();
  for (int o = threadIdx.x; o < p.TB; o += blockDim.x) {
    float acc = p.bcat[o];
    for (int k = 0; k < p.D; ++k) acc = fmaf(p.wcat[o * p.D + k], m[k], acc);
    out[(size_t)row * p.TB + o] = acc;
  }
}

cudaError_t launch_time_bias(const TimeMlpParams& p, const float* t_dev, int nt, float* out, cudaStream_t s) {
  const size_t smem = (size_t)(6 * p.D) * sizeof(float);
  time_bias_kernel<<<nt, 128, smem, s>>>(p, t_dev, out);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) ddpm_step_kernel(StepParams p) {
  ddpm_step_range<false>(p, 0, p.n_elems >> 2, (long long)blockIdx.x * blockDim.x + threadIdx.x,
                         (long long)gridDim.x * blockDim.x);
}

// Stand-alone K4, tiled: a block owns 32 consecutive (sample, frame) rows (32 T floats: a whole number of float4).
// Phase 1 is the float4-vectorised update of ddpm_step_range (same arithmetic, same Philox indexing, so results are
// bit-identical); the new values are kept in shared memory.  Phase 2 writes the fp16 hi/lo operand rows as WHOLE
// 16-byte pieces of full xs_C-channel rows (zero padding included): every 32-byte sector is written completely, where
// the scalar 2-byte stores of the generic path leave 70 of 128 bytes per row dirty and cost L2 a read-modify-write.
constexpr int kK4Rows = 32;

__global__ void __launch_bounds__(256) ddpm_step_tiled_kernel(StepParams p) {
  extern __shared__ __align__(16) float k4_tile[];   // [32][T]
  const int T = p.T, H = p.H, HT = p.H * p.T;
  const long long n_rows = (long long)p.B * H;
  const long long n_tiles = (n_rows + kK4Rows - 1) / kK4Rows;
  const int pieces_per_row = p.xs_C >> 3;
  const CondView cv = resolve_cond(p);
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long r0 = tile * kK4Rows;
    const int rows = (int)min((long long)kK4Rows, n_rows - r0);
    const long long i4_0 = (r0 * T) >> 2;
    const int n4 = (rows * T) >> 2;
    for (int j = threadIdx.x; j < n4; j += blockDim.x) {
      const long long i4 = i4_0 + j;
      const long long e0 = i4 << 2;
      const float4 x4 = reinterpret_cast<const float4*>(p.x)[i4];
      const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
      float ov[4] = {0.f, 0.f, 0.f, 0.f}, zv[4] = {0.f, 0.f, 0.f, 0.f};
      if (p.mode == 0) {
        const float4 t4 = __ldg(reinterpret_cast<const float4*>(p.model_out) + i4);
        ov[0] = t4.x; ov[1] = t4.y; ov[2] = t4.z; ov[3] = t4.w;
      }
      const bool need_z = (p.mode == 0) ? (p.add_noise != 0) : (p.have_start == 0);
      if (need_z) {
        float4 t4;
        if (p.noise) t4 = __ldg(reinterpret_cast<const float4*>(p.noise) + i4);
        else t4 = philox_normal4(p.seed_ptr ? *p.seed_ptr : p.seed, p.rng_offset + (unsigned long long)i4, p.rng_stream);
        zv[0] = t4.x; zv[1] = t4.y; zv[2] = t4.z; zv[3] = t4.w;
      }
      float rv[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (p.mode == 0) {
          float x0 = ov[k];
          if (p.predict_epsilon) x0 = __fsub_rn(__fmul_rn(p.sr, xv[k]), __fmul_rn(p.srm1, ov[k]));
          if (p.clip_denoised) x0 = fminf(fmaxf(x0, -1.f), 1.f);
          const float mean = __fadd_rn(__fmul_rn(p.c1, x0), __fmul_rn(p.c2, xv[k]));
          rv[k] = __fadd_rn(mean, __fmul_rn(p.sigma, zv[k]));
        } else {
          rv[k] = p.have_start ? xv[k] : zv[k];
        }
      }
      // inpaint: HT is a multiple of 4 and so is e0, so the four values belong to one sample
      const long long b = e0 / HT;
      const int rem = (int)(e0 - b * HT);
      if (cv.mask) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (cv.mask[(size_t)b * cv.mask_bstride + rem + k]) rv[k] = cv.value[(size_t)b * cv.value_bstride + rem + k];
      }
      if (cv.poison) rv[0] = rv[1] = rv[2] = rv[3] = __int_as_float(0x7fc00000);
      const float4 r4 = make_float4(rv[0], rv[1], rv[2], rv[3]);
      reinterpret_cast<float4*>(p.x)[i4] = r4;
      if (p.chain) *reinterpret_cast<float4*>(p.chain + ((size_t)b * p.chain_len + p.chain_slot) * HT + rem) = r4;
      reinterpret_cast<float4*>(k4_tile)[j] = r4;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < rows * pieces_per_row; j += blockDim.x) {
      const int r = j / pieces_per_row, q = j - r * pieces_per_row;
      const long long R = r0 + r;
      const long long b = R / H;
      const int h = (int)(R - b * H);
      const size_t row = ((size_t)(b >> 4) * H + h) * kGroupSamples + (b & 15);
      float v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) { const int c = q * 8 + k; v[k] = c < T ? k4_tile[r * T + c] : 0.f; }
      __half2 hi[4], lo[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) split_half2(v[2 * k], v[2 * k + 1], hi[k], lo[k]);
      __half* dst = p.xs + row * p.xs_C + q * 8;
      *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(hi);
      *reinterpret_cast<uint4*>(dst + p.xs_plane_stride) = *reinterpret_cast<const uint4*>(lo);
    }
    __syncthreads();
  }
}

cudaError_t launch_ddpm_step(const StepParams& p, cudaStream_t s) {
  const long long n_rows = (long long)p.B * p.H;
  const bool tiled = (p.xs_C & 7) == 0 && ((p.H * p.T) & 3) == 0 && (p.xs_plane_stride & 7) == 0 &&
                     (reinterpret_cast<uintptr_t>(p.chain) & 15) == 0 && (reinterpret_cast<uintptr_t>(p.xs) & 15) == 0;
  if (tiled) {
    long long blocks = (n_rows + kK4Rows - 1) / kK4Rows;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    const size_t smem = (size_t)kK4Rows * p.T * sizeof(float);
    ddpm_step_tiled_kernel<<<(unsigned)blocks, 256, smem, s>>>(p);
    return cudaGetLastError();
  }
  const long long n4 = p.n_elems >> 2;
  long long blocks = (n4 + 255) / 256;
  if (blocks > 148LL * 16) blocks = 148LL * 16;
  if (blocks < 1) blocks = 1;
  ddpm_step_kernel<<<(unsigned)blocks, 256, 0, s>>>(p);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// layout helpers: public fp32 [B,H,T] -> operand planes ; operand planes -> fp32 [B,C,L] (debug / tests)
// ------------------------------------------------------------------------------------------------
__global__ void pack_input_kernel(const float* __restrict__ x, __half* xs, long long plane_stride, int xs_C, int B,
                                  int H, int T) {
  const long long n = (long long)B * H * T;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / (H * T);
    const int rem = (int)(i - b * H * T);
    const int h = rem / T, c = rem - h * T;
    const size_t row = ((size_t)(b >> 4) * H + h) * kGroupSamples + (b & 15);
    __half hi, lo;
    split_half(x[i], hi, lo);
    xs[row * xs_C + c] = hi;
    xs[plane_stride + row * xs_C + c] = lo;
  }
}
cudaError_t launch_pack_input(const float* x, __half* xs, long long plane_stride, int xs_C, int B, int H, int T,
                              cudaStream_t s) {
  const long long n = (long long)B * H * T;
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  pack_input_kernel<<<(unsigned)blocks, 256, 0, s>>>(x, xs, plane_stride, xs_C, B, H, T);
  return cudaGetLastError();
}

__global__ void unpack_act_kernel(const __half* __restrict__ a, long long plane_stride, int C, int L, int B,
                                  float* __restrict__ out) {
  const long long n = (long long)B * C * L;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / ((long long)C * L);
    const int rem = (int)(i - b * C * L);
    const int c = rem / L, l = rem - c * L;
    const size_t row = ((size_t)(b >> 4) * L + l) * kGroupSamples + (b & 15);
    out[i] = __half2float(a[row * C + c]) + __half2float(a[plane_stride + row * C + c]);
  }
}
cudaError_t launch_unpack_act(const __half* a, long long plane_stride, int C, int L, int B, float* out,
                              cudaStream_t s) {
  const long long n = (long long)B * C * L;
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  unpack_act_kernel<<<(unsigned)blocks, 256, 0, s>>>(a, plane_stride, C, L, B, out);
  return cudaGetLastError();
}

}  // namespace dm
