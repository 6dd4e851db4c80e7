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

cudaError_t launch_ddpm_step(const StepParams& p, cudaStream_t s) {
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
