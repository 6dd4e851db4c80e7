/* dm_b200.h -- C ABI of the B200-native DiffMimic sampling hot path (libdm_b200.so).
 *
 * The reference (wallaronsk/DeepMimic_diffusion_mujoco) has no FFI: its boundary for this path is the Python
 * class API.  Each entry point below names the reference interface it replaces (paths under
 * /root/reference/diffusion/).  All pointers suffixed _dev are CUDA device pointers on the current device;
 * all others are host pointers.  Every call returns 0 on success or a negative dm_status; the message is in
 * dm_last_error().  No exceptions cross this boundary; the caller owns every buffer it passes.
 * There is NO CPU fallback: without an sm_100 device every compute call fails with DM_ERR_DEVICE.
 */
#ifndef DM_B200_H_
#define DM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dm_plan dm_plan_t;
typedef void* dm_stream_t; /* cudaStream_t */

enum dm_status {
  DM_OK = 0,
  DM_ERR_ARG = -1,     /* bad argument / unsupported shape (e.g. H % 8 != 0) */
  DM_ERR_DEVICE = -2,  /* no CUDA device, or device is not sm_100 */
  DM_ERR_CUDA = -3,    /* a CUDA runtime / driver call failed */
  DM_ERR_WEIGHT = -4,  /* a state_dict tensor is missing or has the wrong shape */
  DM_ERR_STATE = -5    /* call order violated (e.g. sampling before dm_plan_set_schedule) */
};

/* One entry of a reference TemporalUnet.state_dict(): name as produced by the reference module
 * (e.g. "downs.0.0.blocks.0.block.0.weight"), fp32 contiguous host data. */
typedef struct {
  const char* name;
  const float* data;
  int ndim;
  int64_t shape[4];
} dm_tensor_t;

/* The five schedule buffers the sampler reads (GaussianDiffusion.__init__, diffuser/models/diffusion.py:64-111),
 * host fp32 arrays of length n_timesteps. */
typedef struct {
  int n_timesteps;
  const float* sqrt_recip_alphas_cumprod;
  const float* sqrt_recipm1_alphas_cumprod;
  const float* posterior_mean_coef1;
  const float* posterior_mean_coef2;
  const float* posterior_log_variance_clipped;
  /* Optional: the timestep value fed to the time embedding at each schedule index (NULL: index i embeds t = i).
   * DiffusionV4.sample (diffuser/models/diffusion_v4.py:100-147) walks t = T-1..1, i.e. index i embeds t = i + 1. */
  const float* time_values;
} dm_schedule_t;

/* Conditioning ("inpaint") spec: x = mask ? value : x, applied before the first step and after every step.
 * Replaces conditioning_fn / apply_conditioning* (diffuser/models/sampling_config.py:6-32; diffusion.py:206,215).
 * mask_dev: uint8 [H*T] (batched = 0) or [B,H*T] (batched = 1); value_dev likewise fp32.  NULL mask = identity. */
typedef struct {
  const uint8_t* mask_dev;
  const float* value_dev;
  int mask_batched;
  int value_batched;
  /* Optional device-resident decision, uint8[4] = {any entry masked, mask is batch-uniform, value is batch-uniform,
   * the function is a pure overwrite}.  When non-NULL, mask_dev / value_dev must be the full [B,H*T] tensors
   * (mask_batched = value_batched = 1) and the KERNEL picks row 0 for batch-uniform tensors, skips the inpaint when
   * nothing is masked, and poisons x with NaN when the traced function was not a pure overwrite -- so a caller's
   * tracer never needs a device->host read-back, and a stale host-side guess can never select the wrong rows. */
  const uint8_t* flags_dev;
} dm_cond_t;

const char* dm_last_error(void);

/* TemporalUnet.__init__ + load_state_dict (diffuser/models/temporal.py:51-113).  Packs the weights into fp16
 * hi/lo UMMA operand tiles, builds TMA descriptors and the per-layer launch list for batch B, horizon H,
 * transition_dim T.  attention=True state dicts (helpers.py:75-131) add four jobs per attention block (LayerNorm stage,
 * to_qkv conv, softmax / context stage, to_out conv) to the same list; dm_sample_loop stays one launch.  The layout is
 * chosen here: one CTA pair per 16-sample group, two pairs per group for small batches at H >= 128 (dm_plan_ctas_per_group),
 * clusters of two groups sharing weight tiles by TMA multicast at H <= 128 when all of them are resident at once. */
int dm_plan_create(const dm_tensor_t* weights, int n_weights, int B, int H, int T, int dim, const int* dim_mults,
                   int n_mults, dm_plan_t** plan);
void dm_plan_free(dm_plan_t* plan);

/* GaussianDiffusion.__init__ schedule buffers (diffusion.py:64-111); also precomputes the per-timestep
 * time-embedding rows (temporal.py:67-72,28-32) for t = 0..n_timesteps-1. */
int dm_plan_set_schedule(dm_plan_t* plan, const dm_schedule_t* sched);

/* TemporalUnet.forward(x, cond, time) (diffuser/models/temporal.py:115-165).  x_dev, out_dev: fp32 [B,H,T];
 * t_dev: fp32 [B] timestep values (cond is ignored by the reference and has no counterpart here). */
int dm_unet_forward(dm_plan_t* plan, const float* x_dev, const float* t_dev, float* out_dev, dm_stream_t stream);

/* default_sample_fn + p_mean_variance + q_posterior + conditioning_fn for ONE timestep t, given the model
 * output (diffusion.py:19-29,149-184,215).  x_dev is updated in place.  noise_dev NULL -> Philox(seed). */
int dm_ddpm_step(dm_plan_t* plan, float* x_dev, const float* model_out_dev, const float* noise_dev, uint64_t seed,
                 int t, int predict_epsilon, int clip_denoised, const dm_cond_t* cond, dm_stream_t stream);

/* GaussianDiffusion.p_sample_loop (diffusion.py:186-228) up to, not including, sort_by_values (:225).
 *   x_dev        fp32 [B,H,T]; if have_start it holds starting_motion on entry; holds the sample on return
 *   noise_stack_dev  NULL (Philox) or fp32 [timesteps+1,B,H,T]: slot 0 = initial draw, slot k = k-th step's draw
 *   timesteps    n_timesteps, or max_timesteps (:210)
 *   chain_dev    NULL or fp32 [B,timesteps+1,H,T] (return_chain, :208,220-227)
 *   fused        1: ONE launch of the persistent cluster kernel (a 2-CTA cluster owns 16 samples and runs every
 *                layer of every step with cluster-local barriers); 0: one launch per layer (debug / A-B check) */
int dm_sample_loop(dm_plan_t* plan, float* x_dev, int have_start, const float* noise_stack_dev, uint64_t seed,
                   int timesteps, int predict_epsilon, int clip_denoised, const dm_cond_t* cond, float* chain_dev,
                   int fused, dm_stream_t stream);

/* Batched humanoid forward kinematics (MuJoCo mj_kinematics semantics on assets/dp_env_v2.xml; replaces
 * sim.forward() at mocap_player.py:67-73) + DeepMimic tracking reward.
 *   qpos_dev [N,H,35]; ref_dev [H,35]; body_pos_dev NULL or [N,H,15,3] (13 bodies + 2 wrists);
 *   reward_dev [N,H]; sums_dev double[8]: sum r, sum r^2, sum pose, vel, end-effector, com terms, count, 0 */
int dm_fk_reward(const float* qpos_dev, const float* ref_dev, int N, int H, float dt, float* body_pos_dev,
                 float* reward_dev, double* sums_dev, dm_stream_t stream);

/* dm_fk_reward without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same arguments, kernels and results as dm_fk_reward; the reference-clip feature scratch lives in a static device array
 * of DM_FK_NOALLOC_MAX_H frames x 92 floats (376 832 bytes, held by every GPU that loads the library until exit), so
 * H > DM_FK_NOALLOC_MAX_H is rejected with DM_ERR_ARG.  The call issues no cudaMalloc / cudaFree / pool operation.  All
 * calls share that array and it keeps the previous call's contents (each call rewrites the rows it uses before reading
 * them): calls that overlap on the GPU -- different streams, unordered graph branches -- race; serialise them. */
#define DM_FK_NOALLOC_MAX_H 1024
int dm_fk_reward_noalloc(const float* qpos_dev, const float* ref_dev, int N, int H, float dt, float* body_pos_dev,
                         float* reward_dev, double* sums_dev, dm_stream_t stream);

/* MotionFID.calculate_activation_statistics moments (metrics/fid_score.py:24-43): sum_dev double[D] += sum_n X,
 * sxx_dev double[D,D] += X^T X over X_dev fp32 [N,D].  Buffers accumulate (zero them first); sums are what
 * ranks all-reduce. */
int dm_fid_moments(const float* X_dev, int N, int D, double* sum_dev, double* sxx_dev, dm_stream_t stream);

/* SiFID statistics (eval.py:59-77 calc_sifid -> slice_motion_sample :15-21 -> fid_score.py:24-43), batched over
 * samples: X_dev fp32 [N,H,F]; windows of `window` frames at offsets 0, step, 2 step, ...; per sample
 * mu_dev [N, D] and sigma_dev [N, D, D] (D = window * F, covariance over the sample's W windows, / (W - 1)). */
int dm_window_moments(const float* X_dev, int N, int H, int F, int window, int step, float* mu_dev, float* sigma_dev,
                      dm_stream_t stream);

/* On-device mocap loader (utils/mocap_v2.py:105-291 + utils/mocap_util.py:31-48): DeepMimic frames, host-parsed JSON
 * "Frames" as double [F,44] on the device -> qpos rows fp32 [F,35] (float64 arithmetic, cast like motion_dataset.py:38). */
int dm_mocap_to_qpos(const double* frames_dev, int F, float* qpos_dev, dm_stream_t stream);
/* Velocity rows of the same loader (utils/mocap_v2.py:196-291 data_vel, calc_rot_vel :155-178): frames_dev double [F,44]
 * -> qvel rows fp32 [F,34] (root linear 3, root angular 3, per joint 3 or 1); with dm_mocap_to_qpos these are the 69
 * features of data_loaders/motion_dataset_v2.py:66-83 (the dim=128 TemporalUnet's inputs, eval.py:189-197). */
int dm_mocap_to_qvel(const double* frames_dev, int F, float* qvel_dev, dm_stream_t stream);
/* MotionDataset tensor (data_loaders/motion_dataset.py:31-42): out_dev fp32 [n,n,35], row i = the clip rolled by i frames. */
int dm_motion_dataset(const float* qpos_dev, int n, float* out_dev, dm_stream_t stream);

/* Multi-GPU sharding: global index of this plan's sample 0.  The in-kernel Philox counters are then
 * those of rows [first_sample, first_sample + B) of one global batch, so N shards reproduce the 1-GPU trajectories. */
int dm_plan_set_sample_offset(dm_plan_t* plan, long long first_sample);

/* Test / debug introspection: the activation written by each layer launch, unpacked to fp32 [B,C,L]. */
int dm_plan_num_layers(const dm_plan_t* plan);
int dm_plan_layer_info(const dm_plan_t* plan, int i, char* name, int name_cap, int* C, int* L);
int dm_plan_read_layer(dm_plan_t* plan, int i, float* out_dev, dm_stream_t stream);
/* Test / debug introspection: CTAs that share one 16-sample group -- 2 (one CTA pair), or 4 when the plan splits the
 * positions of every layer between two pairs (small batches at horizons >= 128; DESIGN.md, "L-split").  0 for NULL. */
int dm_plan_ctas_per_group(const dm_plan_t* plan);
/* Debug: record %globaltimer stamps of the phases of every layer of one (cluster, step) of the fused loop into
 * trace_dev (uint64 [num_layers][16]); NULL switches tracing off. */
int dm_plan_set_trace(dm_plan_t* plan, unsigned long long* trace_dev, int cluster, int step);
/* Kernel launches issued by the last dm_unet_forward / dm_sample_loop call. */
int dm_plan_last_launch_count(const dm_plan_t* plan);
/* Non-zero if device code recorded a protocol failure (barrier timeout code). */
int dm_device_error_code(unsigned int* code);

#ifdef __cplusplus
}
#endif
#endif /* DM_B200_H_ */
