// Host-visible parameter blocks and launchers for the elementwise kernels (K3, K4, layout helpers, K5, K6).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "conv_job.h"

namespace dm {

struct TimeMlpParams {
  int D;             // time embedding width (== unet dim)
  int TB;            // total time-bias width: sum of Cout over all residual blocks
  const float* w1;   // [4D, D]   time_mlp.1
  const float* b1;   // [4D]
  const float* w2;   // [D, 4D]   time_mlp.3
  const float* b2;   // [D]
  const float* wcat; // [TB, D]   every block's time_mlp.1, concatenated in block order
  const float* bcat; // [TB]
};

struct StepParams {
  int mode;            // 0 = denoise step, 1 = init (x = starting motion or noise; then inpaint)
  int t_index;         // timestep of this step (row of the time-bias table)
  float* x;            // [B,H,T] fp32, updated in place
  const float* model_out;
  const float* noise;  // injected noise [B,H,T] or null -> Philox(seed, offset)
  unsigned long long seed, rng_offset;
  const unsigned long long* seed_ptr;   // if set, overrides seed (lets one CUDA graph serve many seeds)
  unsigned int rng_stream;
  int add_noise;       // t > 0
  int have_start;      // init mode: keep x (starting_motion) instead of drawing noise
  int predict_epsilon, clip_denoised;
  float sr, srm1, c1, c2, sigma;
  const unsigned char* mask;   // [H*T] (bstride 0) or [B,H*T]; null = no conditioning
  const float* value;
  long long mask_bstride, value_bstride;
  const unsigned char* cond_flags;   // dm_cond_t::flags_dev (device-side {any, mask uniform, value uniform, pure}) or null
  __half* xs;          // operand planes of the first conv: [2][G][H][16][xs_C]
  long long xs_plane_stride;
  int xs_C;
  float* chain;        // [B, chain_len, H, T] or null
  int chain_len, chain_slot;
  int B, H, T;
  long long n_elems;   // B*H*T (multiple of 4 because H % 8 == 0)
};

// Arguments of the persistent sampling-loop kernel (conv_kernel.cu).  All pointers are device pointers.
// The whole block is a __grid_constant__ kernel parameter (~19 KB; CUDA >= 12.1 allows 32 KB).
struct LoopParams {
  ConvJob jobs[kMaxLoopJobs]; // layer descriptors in execution order (constant bank -> uniform registers)
  const ConvMaps* maps;       // [n_jobs] TMA descriptors (global memory)
  int n_jobs;
  const StepParams* steps;    // [n_steps + 1]: entry 0 = init, entry k = k-th executed denoising step
  int n_steps;
  const float* tb_table;      // [n_timesteps, TB]
  int TB;
  int G, B, H, T;
  int k4_local;               // first and last layer are S = 8: the DDPM update hands over inside the CTA
  unsigned long long* trace;  // debug: [n_jobs][16] globaltimer stamps of one (cluster, step); null in production
  int trace_cluster, trace_step;
};
static_assert(sizeof(LoopParams) <= 32764, "LoopParams is a kernel parameter: 32764 bytes at most (shrink ConvJob or kMaxLoopJobs)");
cudaError_t launch_sample_loop(const LoopParams& lp, size_t smem_bytes, cudaStream_t stream);

cudaError_t launch_time_bias(const TimeMlpParams& p, const float* t_dev, int nt, float* out, cudaStream_t s);
cudaError_t launch_ddpm_step(const StepParams& p, cudaStream_t s);
cudaError_t launch_pack_input(const float* x, __half* xs, long long plane_stride, int xs_C, int B, int H, int T,
                              cudaStream_t s);
cudaError_t launch_unpack_act(const __half* a, long long plane_stride, int C, int L, int B, float* out,
                              cudaStream_t s);

}  // namespace dm
