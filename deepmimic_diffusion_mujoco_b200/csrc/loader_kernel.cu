// On-device mocap loader: DeepMimic frames [F, 44] (duration, root pos 3, root quat 4, twelve
// joints in DeepMimic order as quaternions / hinge angles) -> MuJoCo qpos rows [F, 35], and the MotionDataset tensor
// of all cyclic shifts.  Restates, per frame:
//   utils/mocap_util.py:31-48      align_rotation q' = qL * q * qR (+/-90 degrees about x), align_position (x,y,z)->(x,-z,y)
//   utils/mocap_v2.py:105-153      frame parse in BODY_JOINTS_IN_DP_ORDER
//   utils/mocap_v2.py:180-291      emit in BODY_JOINTS order, quaternion -> Euler 'rxyz' (utils/transformations.py:1031-1097)
//   data_loaders/motion_dataset.py:13-42   row i of the dataset = clip rolled by i frames
// The reference computes in float64 and casts to float32 (motion_dataset.py:38); so does this kernel (one thread per
// frame; the clips are a few hundred frames, precision matters here, not throughput).
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/dm_b200.h"

namespace dm {

int fail(int code, const char* fmt, ...);
int check_device();

struct Qd { double w, x, y, z; };

__device__ __forceinline__ Qd qmul_d(Qd a, Qd b) {
  Qd r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  return r;
}
__device__ __forceinline__ Qd align_rotation(const double* q) {
  const double r = 0.70710678118654752440;
  const Qd left = {r, r, 0.0, 0.0}, right = {r, -r, 0.0, 0.0};
  const Qd in = {q[0], q[1], q[2], q[3]};
  return qmul_d(qmul_d(left, in), right);
}
// intrinsic x-y-z Euler angles of quaternion (w, x, y, z): transformations.euler_from_quaternion(axes='rxyz')
__device__ __forceinline__ void euler_rxyz(Qd q, double* out) {
  const double n = q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z;
  const double s = sqrt(2.0 / n);
  const double w = q.w * s, x = q.x * s, y = q.y * s, z = q.z * s;
  const double m00 = 1.0 - y * y - z * z, m01 = x * y - z * w, m02 = x * z + y * w;
  const double m10 = x * y + z * w, m11 = 1.0 - x * x - z * z;
  const double m12 = y * z - x * w, m22 = 1.0 - x * x - y * y;
  const double cy = sqrt(m22 * m22 + m12 * m12);
  double ax, az;
  if (cy > 2.220446049250313e-16 * 4.0) { ax = atan2(m01, m00); az = atan2(m12, m22); }
  else { ax = atan2(-m10, m11); az = 0.0; }
  const double ay = atan2(-m02, cy);
  out[0] = -az; out[1] = -ay; out[2] = -ax;
}

// DeepMimic order: chest neck r_hip r_knee r_ankle r_shoulder r_elbow l_hip l_knee l_ankle l_shoulder l_elbow
// qpos order:      chest neck r_shoulder r_elbow l_shoulder l_elbow r_hip r_knee r_ankle l_hip l_knee l_ankle
__global__ void mocap_to_qpos_kernel(const double* __restrict__ frames, int F, float* __restrict__ qpos) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  const double* fr = frames + (size_t)f * 44;
  float* o = qpos + (size_t)f * 35;
  o[0] = (float)fr[1]; o[1] = (float)(-fr[3]); o[2] = (float)fr[2];
  const Qd rq = align_rotation(fr + 4);
  o[3] = (float)rq.w; o[4] = (float)rq.x; o[5] = (float)rq.y; o[6] = (float)rq.z;
  // offset of each joint inside the DeepMimic frame and inside qpos; width 4 = quaternion, 1 = hinge
  const int dp_off[12] = {8, 12, 16, 20, 21, 25, 29, 30, 34, 35, 39, 43};
  const int dp_hinge[12] = {0, 0, 0, 1, 0, 0, 1, 0, 1, 0, 0, 1};
  const int to_qpos[12] = {7, 10, 21, 24, 25, 13, 16, 28, 31, 32, 17, 20};
#pragma unroll
  for (int j = 0; j < 12; ++j) {
    if (dp_hinge[j]) {
      o[to_qpos[j]] = (float)fr[dp_off[j]];
    } else {
      double e[3];
      euler_rxyz(align_rotation(fr + dp_off[j]), e);
      o[to_qpos[j]] = (float)e[0]; o[to_qpos[j] + 1] = (float)e[1]; o[to_qpos[j] + 2] = (float)e[2];
    }
  }
}

// utils/mocap_v2.py:155-178 calc_rot_vel(seg_0 = frame k, seg_1 = frame k-1, dura): q_diff = conj(q_0) * q_1, velocity =
// angle / dura * axis with pyquaternion 0.9.9's .axis / .angle: normalise unless |1 - |q|^2| < 1e-14; axis = vector /
// |vector| (zeros below 1e-17); angle = 2 atan2(|vector|, scalar) wrapped into (-pi, pi].
__device__ __forceinline__ void rot_vel(Qd q_now, Qd q_prev, double dura, float* out) {
  const Qd conj = {q_now.w, -q_now.x, -q_now.y, -q_now.z};
  Qd d = qmul_d(conj, q_prev);
  const double ss = d.w * d.w + d.x * d.x + d.y * d.y + d.z * d.z;
  if (!(fabs(1.0 - ss) < 1e-14)) {
    const double n = sqrt(ss);
    if (n > 0.0) { d.w /= n; d.x /= n; d.y /= n; d.z /= n; }
  }
  const double nv = sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
  double ax = 0.0, ay = 0.0, az = 0.0;
  if (!(nv < 1e-17)) { ax = d.x / nv; ay = d.y / nv; az = d.z / nv; }
  const double pi = 3.14159265358979323846;
  double ang = fmod(2.0 * atan2(nv, d.w) + pi, 2.0 * pi) - pi;
  if (ang == -pi) ang = pi;
  const double w = ang / dura;
  out[0] = (float)(w * ax); out[1] = (float)(w * ay); out[2] = (float)(w * az);
}

// data_vel rows (utils/mocap_v2.py:196-291): [root linear 3 | root angular 3 | joints in qpos order: 3 or 1] = 34 values.
// Row 0 is zero; row k differences frame k against frame k-1 over durations[k-1] (0.0167 when zero).
__global__ void mocap_to_qvel_kernel(const double* __restrict__ frames, int F, float* __restrict__ qvel) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  float* o = qvel + (size_t)f * 34;
  if (f == 0) {
    for (int i = 0; i < 34; ++i) o[i] = 0.f;
    return;
  }
  const double* fr = frames + (size_t)f * 44;
  const double* pv = fr - 44;
  double dura = pv[0];
  if (dura == 0.0) dura = 0.0167;
  o[0] = (float)((fr[1] - pv[1]) * 1.0 / dura);
  o[1] = (float)(((-fr[3]) - (-pv[3])) * 1.0 / dura);
  o[2] = (float)((fr[2] - pv[2]) * 1.0 / dura);
  rot_vel(align_rotation(fr + 4), align_rotation(pv + 4), dura, o + 3);
  const int dp_off[12] = {8, 12, 16, 20, 21, 25, 29, 30, 34, 35, 39, 43};
  const int dp_hinge[12] = {0, 0, 0, 1, 0, 0, 1, 0, 1, 0, 0, 1};
  // offsets inside the velocity row: qpos offset minus 1 (the root quaternion has 4 values, its velocity 3)
  const int to_qvel[12] = {6, 9, 20, 23, 24, 12, 15, 27, 30, 31, 16, 19};
#pragma unroll
  for (int j = 0; j < 12; ++j) {
    if (dp_hinge[j]) o[to_qvel[j]] = (float)((fr[dp_off[j]] - pv[dp_off[j]]) * 1.0 / dura);
    else rot_vel(align_rotation(fr + dp_off[j]), align_rotation(pv + dp_off[j]), dura, o + to_qvel[j]);
  }
}

// dataset[i][j][:] = qpos[(i + j) % n][:]   (motion_dataset.py:31-37: cat(cfg[i:], cfg[:i]))
__global__ void motion_dataset_kernel(const float* __restrict__ qpos, int n, float* __restrict__ out) {
  const long long total = (long long)n * n * 35;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(e % 35);
    const long long row = e / 35;
    const int j = (int)(row % n), i = (int)(row / n);
    out[e] = qpos[(size_t)((i + j) % n) * 35 + c];
  }
}

}  // namespace dm

using namespace dm;

extern "C" int dm_mocap_to_qpos(const double* frames_dev, int F, float* qpos_dev, dm_stream_t st) {
  if (!frames_dev || !qpos_dev) return fail(DM_ERR_ARG, "null argument");
  if (F < 1) return fail(DM_ERR_ARG, "need at least one frame");
  int rc = check_device();
  if (rc) return rc;
  mocap_to_qpos_kernel<<<(F + 63) / 64, 64, 0, static_cast<cudaStream_t>(st)>>>(frames_dev, F, qpos_dev);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "mocap_to_qpos launch: %s", cudaGetErrorString(e));
  return DM_OK;
}

extern "C" int dm_mocap_to_qvel(const double* frames_dev, int F, float* qvel_dev, dm_stream_t st) {
  if (!frames_dev || !qvel_dev) return fail(DM_ERR_ARG, "null argument");
  if (F < 1) return fail(DM_ERR_ARG, "need at least one frame");
  int rc = check_device();
  if (rc) return rc;
  mocap_to_qvel_kernel<<<(F + 63) / 64, 64, 0, static_cast<cudaStream_t>(st)>>>(frames_dev, F, qvel_dev);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "mocap_to_qvel launch: %s", cudaGetErrorString(e));
  return DM_OK;
}

extern "C" int dm_motion_dataset(const float* qpos_dev, int n, float* out_dev, dm_stream_t st) {
  if (!qpos_dev || !out_dev) return fail(DM_ERR_ARG, "null argument");
  if (n < 1) return fail(DM_ERR_ARG, "need at least one frame");
  int rc = check_device();
  if (rc) return rc;
  const long long total = (long long)n * n * 35;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  motion_dataset_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(st)>>>(qpos_dev, n, out_dev);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "motion_dataset launch: %s", cudaGetErrorString(e));
  return DM_OK;
}
