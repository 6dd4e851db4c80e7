// Residual(PreNorm(LinearAttention)) launch descriptor (reference: diffuser/models/helpers.py:75-131).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stddef.h>

namespace dm {

struct AttnParams {
  const __half* in;        // activation planes [2][G][L][16][C]
  long long in_plane;      // elements per plane
  __half* out;
  long long out_plane;
  int C, L, NS;            // channels, positions, samples per CTA (1, 2 or 4)
  int heads;               // dim_head is 32 (helpers.py:111)
  float eps, scale;
  const float* ln_g;       // [C]   PreNorm.norm.g
  const float* ln_b;       // [C]   PreNorm.norm.b
  const float* wqkv_t;     // [C][3*heads*32]   to_qkv.weight transposed
  const float* wout_t;     // [heads*32][C]     to_out.weight transposed
  const float* bout;       // [C]
};

size_t attention_smem_bytes(int C, int L, int NS);
int attention_pick_ns(int C, int L);   // 0: the tile does not fit in shared memory
cudaError_t launch_linear_attention(const AttnParams& p, int n_groups, cudaStream_t s);

}  // namespace dm
