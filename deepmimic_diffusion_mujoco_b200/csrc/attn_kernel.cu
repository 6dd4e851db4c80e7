// Per-layer launch path of the attention stages (the fused loop runs them inside sample_loop_kernel, conv_kernel.cu).
#include "attn_kernel.h"

#include <mutex>

#include "attn_stages.cuh"

namespace dm {

__global__ void __launch_bounds__(512, 1) attention_stage_kernel(const __grid_constant__ ConvJob job) {
  extern __shared__ __align__(16) float attn_sm[];
  const int g = blockIdx.x >> 1, b0 = (blockIdx.x & 1) * 8;
  if (job.kind == JOB_ATTN_NORM) attn_norm_stage(job, g, b0, threadIdx.x);
  else attn_core_stage(job, g, b0, threadIdx.x, attn_sm);
}

cudaError_t launch_attention_stage(const ConvJob& job, int n_groups, cudaStream_t s) {
  static std::mutex mu;
  static bool done[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!(dev >= 0 && dev < 64 && done[dev])) {
      e = cudaFuncSetAttribute(attention_stage_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCoreSmemBytes);
      if (e != cudaSuccess) return e;
      if (dev >= 0 && dev < 64) done[dev] = true;
    }
  }
  const size_t smem = job.kind == JOB_ATTN_CORE ? kCoreSmemBytes : 0;
  attention_stage_kernel<<<2 * n_groups, 512, smem, s>>>(job);
  return cudaGetLastError();
}

}  // namespace dm
