// Per-layer launches of the two CUDA-core stages of Residual(PreNorm(LinearAttention)) (attn_stages.cuh; reference:
// diffuser/models/helpers.py:75-131).  The fused sampling loop calls the same device functions in place.
#pragma once
#include <cuda_runtime.h>

#include "conv_job.h"

namespace dm {

// job.kind: JOB_ATTN_NORM or JOB_ATTN_CORE.  One CTA per 8 samples (2 per 16-sample group), 512 threads.
cudaError_t launch_attention_stage(const ConvJob& job, int n_groups, cudaStream_t s);

}  // namespace dm
