// Host-side plan: packed weights, activation workspace, TMA descriptors and the per-layer launch list of one
// TemporalUnet (diffuser/models/temporal.py:49-165) at a fixed (B, H, T, dim, dim_mults).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <map>
#include <string>
#include <vector>

#include "attn_kernel.h"
#include "conv_job.h"
#include "step_kernels.h"

namespace dm {

struct Act {
  __half* p = nullptr;   // [2][G][L][16][C]
  int C = 0, L = 0;
  long long plane = 0;   // elements per plane
};

struct Layer {
  int kind = 0;          // == job.kind: JOB_CONV (job, maps), or an attention stage (job only; attn_stages.cuh)
  bool ls_padded = false;   // L-split: the half is not a whole number of MMA tiles (no resident link from or to this layer)
  bool mc_ok = false;    // quarter-tile weight maps exist (weight multicast)
  bool hidden = false;   // intermediate of an attention block: dm_plan_layer_info reports no output for it
  ConvJob job;
  ConvMaps maps;
  int slices = 1;
  size_t smem = 0;
  int tb_off = -1;       // offset into a time-bias row, or -1
  int out_act = -1;      // index into acts (for dm_plan_read_layer), -1 if fp32-only
  int aux_act = -1;      // output of the fused auxiliary 1x1 conv, -1 if none
  int src_acts[2] = {-1, -1};   // activations this layer reads as conv input ...
  int res_act = -1;             // ... and as residual (consumer bookkeeping for ConvJob::no_store)
  std::string name;
};

// Device tables of one schedule (dm_plan_set_schedule): kept per content hash, so that alternating schedules on one
// plan (GaussianDiffusion and DiffusionV4 over the same U-Net, eval.py-style loops) neither re-runs K3 nor leaks.
struct ScheduleTables {
  unsigned long long hash = 0;
  int n = 0;
  float* t_table = nullptr;    // [n] timestep value embedded at each schedule index
  float* tb_table = nullptr;   // [n, TB] time-bias rows
  std::vector<float> sr, srm1, c1, c2, logvar;
};

struct HostTensor {
  const float* data;
  std::vector<int64_t> shape;
};

}  // namespace dm

struct dm_plan {
  int B = 0, H = 0, T = 0, dim = 0, G = 0, device = 0;
  std::vector<int> mults;
  std::vector<void*> allocs;
  std::vector<dm::Act> acts;
  std::vector<dm::Layer> layers;
  std::vector<dm::ConvJob> loop_jobs;   // the fused loop's copy of the descriptors: activations alias a small shared workspace
  size_t loop_workspace_bytes = 0, layer_workspace_bytes = 0;   // activation bytes the fused loop / the per-layer path touch per step
  int xs_act = -1;
  bool mcast = false;              // fused loop: clusters of four = two groups sharing weight tiles by TMA multicast
  bool lsplit = false;             // L-split (conv_job.h): two CTA pairs per 16-sample group, half of the positions each
  bool has_attention = false;      // attention=True: per-layer launches only (the fused loop kernel has no attention stage)
  float* model_out = nullptr;      // fp32 [B,H,T] scratch for the sampling loop
  dm::TimeMlpParams tmlp{};
  int TB = 0;
  float* tb_table = nullptr;       // [n_timesteps, TB]  (points into the active entry of `schedules`)
  std::vector<dm::ScheduleTables> schedules;   // small cache, least recently used entry is replaced
  int steps_capacity = 0;          // entries allocated in steps_dev / steps_host
  float* tb_rows = nullptr;        // [G*16, TB] per-sample rows for dm_unet_forward
  float* t_table = nullptr;        // [n_timesteps] 0..n-1
  int n_timesteps = 0;
  std::vector<float> sr, srm1, c1, c2, logvar;
  int last_launches = 0;
  unsigned long long rng_calls = 0;
  long long sample_offset = 0;     // global index of this plan's sample 0 (multi-GPU shards draw the noise of their global rows)
  unsigned long long* seed_dev = nullptr;
  // persistent sampling-loop kernel arguments
  dm::ConvMaps* maps_dev = nullptr;
  size_t loop_smem = 0;
  dm::StepParams* steps_dev = nullptr;
  dm::StepParams* steps_host = nullptr;   // pinned staging
  cudaEvent_t steps_event = nullptr;
  bool steps_pending = false;
  unsigned long long* trace_dev = nullptr;   // debug phase trace (dm_plan_set_trace)
  int trace_cluster = 0, trace_step = 0;
};
