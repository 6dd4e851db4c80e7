// K5: batched humanoid forward kinematics + DeepMimic tracking reward; K6: motion-FID moment accumulation.
//
// FK follows MuJoCo mj_kinematics on /root/reference/diffusion/assets/dp_env_v2.xml:20-107 (the reference
// only reaches FK through mujoco_py sim.forward(), mocap_player.py:67-73): free root = qpos[0:7]; a child body
// starts at parent_pos + R_parent * body_offset with the parent's orientation; each hinge, in document order,
// rotates the frame about its body-frame axis through its anchor.
// Reward: SURVEY.md 8(a) a15 (this repo's spec; weights from utils/mocap_util.py:26-29).
// FID moments: metrics/fid_score.py:24-43 restated as streaming sums.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/dm_b200.h"

namespace dm {

int fail(int code, const char* fmt, ...);
int check_device();

struct V3 { float x, y, z; };
struct Q4 { float w, x, y, z; };

__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ Q4 qmul(Q4 a, Q4 b) {
  Q4 r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  return r;
}
__device__ __forceinline__ Q4 qconj(Q4 a) { Q4 r; r.w = a.w; r.x = -a.x; r.y = -a.y; r.z = -a.z; return r; }
__device__ __forceinline__ V3 qrot(Q4 q, V3 v) {
  // v + 2 w (u x v) + 2 u x (u x v)
  const float cx = q.y * v.z - q.z * v.y, cy = q.z * v.x - q.x * v.z, cz = q.x * v.y - q.y * v.x;
  const float dx = q.y * cz - q.z * cy, dy = q.z * cx - q.x * cz, dz = q.x * cy - q.y * cx;
  return v3(v.x + 2.f * (q.w * cx + dx), v.y + 2.f * (q.w * cy + dy), v.z + 2.f * (q.w * cz + dz));
}
__device__ __forceinline__ Q4 qaxis(int axis, float sign, float ang) {
  float s, c;
  sincosf(0.5f * ang, &s, &c);
  Q4 r; r.w = c; r.x = 0.f; r.y = 0.f; r.z = 0.f;
  if (axis == 0) r.x = sign * s; else if (axis == 1) r.y = sign * s; else r.z = sign * s;
  return r;
}
__device__ __forceinline__ float qangle(Q4 q) {   // rotation angle in [0, pi]
  const float vn = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z);
  return 2.f * atan2f(vn, fabsf(q.w));
}
__device__ __forceinline__ V3 qlogvel(Q4 q, float inv_dt) {   // axis * angle / dt (mocap_v2.py:155-178)
  const float vn = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z);
  if (vn < 1e-12f) return v3(0.f, 0.f, 0.f);
  float w = q.w, sx = 1.f;
  if (w < 0.f) { w = -w; sx = -1.f; }
  const float k = 2.f * atan2f(vn, w) * inv_dt * sx / vn;
  return v3(q.x * k, q.y * k, q.z * k);
}

// rotate frame (pos, q) by a hinge about body-frame `axis` through body-frame anchor (ax, 0, 0)
__device__ __forceinline__ void hinge(V3& pos, Q4& q, int axis, float sign, float anchor_x, float ang) {
  if (anchor_x != 0.f) {
    const V3 a = vadd(pos, qrot(q, v3(anchor_x, 0.f, 0.f)));
    q = qmul(q, qaxis(axis, sign, ang));
    pos = vsub(a, qrot(q, v3(anchor_x, 0.f, 0.f)));
  } else {
    q = qmul(q, qaxis(axis, sign, ang));
  }
}

struct Feat {
  Q4 rootq;
  V3 rootw;
  Q4 jq[8];      // local rotations of the eight 3-DoF joints (chest, neck, r/l shoulder, r/l hip, r/l ankle order below)
  float ja[4];   // elbow / knee angles
  float rate[28];
  V3 ee[4];      // r_ankle, l_ankle, r_wrist, l_wrist relative to the root position
  V3 com;
};
constexpr int kFeatFloats = sizeof(Feat) / sizeof(float);

// body order of body_pos output == BODY_DEFS (utils/mocap_util.py:17-19)
enum { B_ROOT, B_CHEST, B_NECK, B_RHIP, B_RKNEE, B_RANKLE, B_RSHO, B_RELB, B_RWRIST, B_LHIP, B_LKNEE, B_LANKLE, B_LSHO,
       B_LELB, B_LWRIST, B_COUNT };

__device__ __forceinline__ void ball(V3& pos, Q4& q, const float* a, float anchor_y_hinge_x) {
  hinge(pos, q, 0, 1.f, 0.f, a[0]);
  hinge(pos, q, 1, 1.f, anchor_y_hinge_x, a[1]);
  hinge(pos, q, 2, 1.f, 0.f, a[2]);
}
__device__ __forceinline__ Q4 euler_q(const float* a) {
  return qmul(qmul(qaxis(0, 1.f, a[0]), qaxis(1, 1.f, a[1])), qaxis(2, 1.f, a[2]));
}

// q: 35 qpos of this frame; qp: previous frame (== q for frame 0 -> zero velocity, mocap_v2.py:201-205)
__device__ void frame_features(const float* q, const float* qp, float inv_dt, Feat& f, V3* bp) {
  Q4 rq; rq.w = q[3]; rq.x = q[4]; rq.y = q[5]; rq.z = q[6];
  float n = sqrtf(rq.w * rq.w + rq.x * rq.x + rq.y * rq.y + rq.z * rq.z);
  if (n < 1e-10f) { rq.w = 1.f; rq.x = rq.y = rq.z = 0.f; } else { n = 1.f / n; rq.w *= n; rq.x *= n; rq.y *= n; rq.z *= n; }
  Q4 pq; pq.w = qp[3]; pq.x = qp[4]; pq.y = qp[5]; pq.z = qp[6];
  n = sqrtf(pq.w * pq.w + pq.x * pq.x + pq.y * pq.y + pq.z * pq.z);
  if (n < 1e-10f) { pq.w = 1.f; pq.x = pq.y = pq.z = 0.f; } else { n = 1.f / n; pq.w *= n; pq.x *= n; pq.y *= n; pq.z *= n; }
  f.rootq = rq;
  f.rootw = qlogvel(qmul(qconj(pq), rq), inv_dt);
  const V3 rp = v3(q[0], q[1], q[2]);
  bp[B_ROOT] = rp;

  V3 p; Q4 o;
  // chest (qpos 7-9), neck (10-12)
  p = vadd(rp, qrot(rq, v3(0.f, 0.f, 0.236151f))); o = rq;
  ball(p, o, q + 7, 0.f);
  const V3 chest_p = p; const Q4 chest_q = o;
  bp[B_CHEST] = p;
  p = vadd(chest_p, qrot(chest_q, v3(0.f, 0.f, 0.223894f))); o = chest_q;
  ball(p, o, q + 10, 0.f);
  bp[B_NECK] = p;
  const V3 neck_p = p; const Q4 neck_q = o;
  // arms: shoulder (13-15 / 17-19), elbow (16 / 20): axis (0,-1,0), anchor (0.028,0,0); wrist geom at (0,0,-0.258947)
  V3 sh_p[2], el_p[2], wr_p[2]; Q4 sh_q[2], el_q[2];
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    const float ysign = s == 0 ? -1.f : 1.f;
    const int qi = s == 0 ? 13 : 17;
    p = vadd(chest_p, qrot(chest_q, v3(-0.02405f, ysign * 0.18311f, 0.2435f))); o = chest_q;
    ball(p, o, q + qi, 0.f);
    sh_p[s] = p; sh_q[s] = o;
    p = vadd(p, qrot(o, v3(0.f, 0.f, -0.274788f)));
    hinge(p, o, 1, -1.f, 0.028f, q[qi + 3]);
    el_p[s] = p; el_q[s] = o;
    wr_p[s] = vadd(p, qrot(o, v3(0.f, 0.f, -0.258947f)));
  }
  bp[B_RSHO] = sh_p[0]; bp[B_RELB] = el_p[0]; bp[B_RWRIST] = wr_p[0];
  bp[B_LSHO] = sh_p[1]; bp[B_LELB] = el_p[1]; bp[B_LWRIST] = wr_p[1];
  // legs: hip (21-23 / 28-30; y hinge anchored at (0.02,0,0)), knee (24 / 31; axis (0,-1,0), anchor (-0.035,0,0)),
  // ankle (25-27 / 32-34; y hinge anchored at (-0.02,0,0))
  V3 hip_p[2], kn_p[2], an_p[2]; Q4 hip_q[2], kn_q[2], an_q[2];
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    const float ysign = s == 0 ? -1.f : 1.f;
    const int qi = s == 0 ? 21 : 28;
    p = vadd(rp, qrot(rq, v3(0.f, ysign * 0.094887f, -0.05f))); o = rq;
    ball(p, o, q + qi, 0.02f);
    hip_p[s] = p; hip_q[s] = o;
    p = vadd(p, qrot(o, v3(0.f, 0.f, -0.421546f)));
    hinge(p, o, 1, -1.f, -0.035f, q[qi + 3]);
    kn_p[s] = p; kn_q[s] = o;
    p = vadd(p, qrot(o, v3(0.f, 0.f, -0.40987f)));
    ball(p, o, q + qi + 4, -0.02f);
    an_p[s] = p; an_q[s] = o;
  }
  bp[B_RHIP] = hip_p[0]; bp[B_RKNEE] = kn_p[0]; bp[B_RANKLE] = an_p[0];
  bp[B_LHIP] = hip_p[1]; bp[B_LKNEE] = kn_p[1]; bp[B_LANKLE] = an_p[1];

  // local joint rotations in qpos (BODY_JOINTS) order
  f.jq[0] = euler_q(q + 7);  f.jq[1] = euler_q(q + 10); f.jq[2] = euler_q(q + 13); f.jq[3] = euler_q(q + 17);
  f.jq[4] = euler_q(q + 21); f.jq[5] = euler_q(q + 25); f.jq[6] = euler_q(q + 28); f.jq[7] = euler_q(q + 32);
  f.ja[0] = q[16]; f.ja[1] = q[20]; f.ja[2] = q[24]; f.ja[3] = q[31];
#pragma unroll
  for (int d = 0; d < 28; ++d) f.rate[d] = (q[7 + d] - qp[7 + d]) * inv_dt;
  f.ee[0] = vsub(an_p[0], rp); f.ee[1] = vsub(an_p[1], rp);
  f.ee[2] = vsub(wr_p[0], rp); f.ee[3] = vsub(wr_p[1], rp);

  // centre of mass from the XML geoms (mass, body-local centre)
  V3 c = v3(0.f, 0.f, 0.f);
  auto acc = [&](float m, V3 bpz, Q4 bq, V3 local) {
    const V3 w = vadd(bpz, qrot(bq, local));
    c.x += m * w.x; c.y += m * w.y; c.z += m * w.z;
  };
  acc(2.5f, rp, rq, v3(0.f, 0.f, 0.07f));
  acc(3.0f, chest_p, chest_q, v3(0.f, 0.f, 0.12f));
  acc(2.0f, neck_p, neck_q, v3(0.f, 0.f, 0.175f));
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    acc(1.5f, sh_p[s], sh_q[s], v3(0.f, 0.f, -0.14f));
    acc(1.0f, el_p[s], el_q[s], v3(0.f, 0.f, -0.12f));
    acc(0.5f, el_p[s], el_q[s], v3(0.f, 0.f, -0.258947f));
    acc(4.5f, hip_p[s], hip_q[s], v3(0.f, 0.f, -0.21f));
    acc(3.0f, kn_p[s], kn_q[s], v3(0.f, 0.f, -0.2f));
    acc(1.0f, an_p[s], an_q[s], v3(0.045f, 0.f, -0.0425f));
  }
  const float inv_m = 1.f / 30.5f;
  f.com = v3(c.x * inv_m, c.y * inv_m, c.z * inv_m);
}

__global__ void ref_features_kernel(const float* __restrict__ ref, int H, float inv_dt, float* __restrict__ feat) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= H) return;
  float q[35], qp[35];
  for (int i = 0; i < 35; ++i) { q[i] = ref[h * 35 + i]; qp[i] = ref[(h > 0 ? h - 1 : 0) * 35 + i]; }
  Feat f;
  V3 bp[B_COUNT];
  frame_features(q, qp, inv_dt, f, bp);
  const float* src = reinterpret_cast<const float*>(&f);
  for (int i = 0; i < kFeatFloats; ++i) feat[(size_t)h * kFeatFloats + i] = src[i];
}

constexpr int kFkThreads = 128;

__global__ void __launch_bounds__(kFkThreads) fk_reward_kernel(const float* __restrict__ qpos, const float* __restrict__ feat,
                                                               long long n_frames, int H, float inv_dt,
                                                               float* __restrict__ body_pos, float* __restrict__ reward,
                                                               double* __restrict__ sums) {
  __shared__ float sq[(kFkThreads + 1) * 35];
  __shared__ float red[6][kFkThreads / 32];
  const long long f0 = (long long)blockIdx.x * kFkThreads;
  const long long nf = min((long long)kFkThreads, n_frames - f0);
  // coalesced stage of frames [f0-1, f0+nf) into shared memory (row 0 = the frame before f0)
  {
    const long long g0 = (f0 > 0 ? f0 - 1 : 0) * 35;
    const long long g1 = (f0 + nf) * 35;
    const int skip = f0 > 0 ? 0 : 35;
    for (long long i = g0 + threadIdx.x; i < g1; i += kFkThreads) sq[skip + (int)(i - g0)] = __ldg(qpos + i);
  }
  __syncthreads();
  float r = 0.f, tp = 0.f, tv = 0.f, te = 0.f, tc = 0.f;
  const long long fi = f0 + threadIdx.x;
  const bool active = threadIdx.x < nf;
  if (active) {
    const int h = (int)(fi % H);
    float q[35], qp[35];
#pragma unroll
    for (int i = 0; i < 35; ++i) q[i] = sq[(threadIdx.x + 1) * 35 + i];
#pragma unroll
    for (int i = 0; i < 35; ++i) qp[i] = h > 0 ? sq[threadIdx.x * 35 + i] : q[i];
    Feat f;
    V3 bp[B_COUNT];
    frame_features(q, qp, inv_dt, f, bp);
    const Feat& g = *reinterpret_cast<const Feat*>(feat + (size_t)h * kFeatFloats);
    const float w3[8] = {0.5f, 0.3f, 0.3f, 0.3f, 0.5f, 0.2f, 0.5f, 0.2f};   // chest neck rsho lsho rhip rank lhip lank
    const float w1[4] = {0.2f, 0.2f, 0.3f, 0.3f};                           // relbow lelbow rknee lknee
    // pose
    float a = qangle(qmul(qconj(g.rootq), f.rootq));
    float pose = a * a;
#pragma unroll
    for (int k = 0; k < 8; ++k) { a = qangle(qmul(qconj(g.jq[k]), f.jq[k])); pose += w3[k] * a * a; }
#pragma unroll
    for (int k = 0; k < 4; ++k) { a = f.ja[k] - g.ja[k]; pose += w1[k] * a * a; }
    // velocity: rate index d <-> qpos 7+d
    V3 dw = vsub(f.rootw, g.rootw);
    float vel = dw.x * dw.x + dw.y * dw.y + dw.z * dw.z;
    const float wr[28] = {0.5f, 0.5f, 0.5f, 0.3f, 0.3f, 0.3f, 0.3f, 0.3f, 0.3f, 0.2f, 0.3f, 0.3f, 0.3f, 0.2f,
                          0.5f, 0.5f, 0.5f, 0.3f, 0.2f, 0.2f, 0.2f, 0.5f, 0.5f, 0.5f, 0.3f, 0.2f, 0.2f, 0.2f};
#pragma unroll
    for (int d = 0; d < 28; ++d) { a = f.rate[d] - g.rate[d]; vel += wr[d] * a * a; }
    float ee = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) { V3 d = vsub(f.ee[k], g.ee[k]); ee += d.x * d.x + d.y * d.y + d.z * d.z; }
    V3 dc = vsub(f.com, g.com);
    const float com = dc.x * dc.x + dc.y * dc.y + dc.z * dc.z;
    tp = expf(-2.f * pose); tv = expf(-0.1f * vel); te = expf(-40.f * ee); tc = expf(-10.f * com);
    r = 0.65f * tp + 0.1f * tv + 0.15f * te + 0.1f * tc;
    reward[fi] = r;
    if (body_pos) {
      float* o = body_pos + (size_t)fi * B_COUNT * 3;
#pragma unroll
      for (int k = 0; k < B_COUNT; ++k) { o[3 * k] = bp[k].x; o[3 * k + 1] = bp[k].y; o[3 * k + 2] = bp[k].z; }
    }
  }
  // block reduction -> one double atomic per statistic per block
  float vals[6] = {r, r * r, tp, tv, te, tc};
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float v = vals[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    double s = 0.0;
    for (int w = 0; w < kFkThreads / 32; ++w) s += (double)red[threadIdx.x][w];
    atomicAdd(&sums[threadIdx.x], s);
  }
  if (threadIdx.x == 6) atomicAdd(&sums[6], (double)nf);
}

// ------------------------------------------------------------------------------------------------
// K6: sum[D] += sum_n x ; sxx[D,D] += X^T X.  64x64 output tile per CTA (upper-triangular tiles, mirrored),
// rows split over gridDim.z, fp32 register-tiled accumulation per split, fp64 atomics across splits.
// ------------------------------------------------------------------------------------------------
constexpr int kFidTile = 64, kFidK = 32;

__global__ void __launch_bounds__(256) fid_moments_kernel(const float* __restrict__ X, int N, int D, int rows_per_split,
                                                          double* __restrict__ sum, double* __restrict__ sxx) {
  // map linear block index to an upper-triangular tile (ti <= tj)
  const int nt = (D + kFidTile - 1) / kFidTile;
  int ti = 0, rem = blockIdx.x;
  while (rem >= nt - ti) { rem -= nt - ti; ++ti; }
  const int tj = ti + rem;
  const int i0 = ti * kFidTile, j0 = tj * kFidTile;
  const int r0 = blockIdx.z * rows_per_split;
  const int r1 = min(N, r0 + rows_per_split);
  __shared__ float sa[kFidK][kFidTile + 4], sb[kFidK][kFidTile + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
  float colsum = 0.f;   // threads 0..63 of diagonal tiles accumulate the column sums
  for (int r = r0; r < r1; r += kFidK) {
    for (int e = threadIdx.x; e < kFidK * kFidTile; e += 256) {
      const int k = e / kFidTile, c = e % kFidTile;
      const int row = r + k;
      sa[k][c] = (row < r1 && i0 + c < D) ? __ldg(X + (size_t)row * D + i0 + c) : 0.f;
      sb[k][c] = (row < r1 && j0 + c < D) ? __ldg(X + (size_t)row * D + j0 + c) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < kFidK; ++k) {
      const float4 av = *reinterpret_cast<const float4*>(&sa[k][ty * 4]);
      const float4 bv = *reinterpret_cast<const float4*>(&sb[k][tx * 4]);
      const float a[4] = {av.x, av.y, av.z, av.w}, b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[p][q] = fmaf(a[p], b[q], acc[p][q]);
    }
    if (ti == tj && threadIdx.x < kFidTile) {
      for (int k = 0; k < kFidK; ++k) colsum += sa[k][threadIdx.x];
    }
    __syncthreads();
  }
#pragma unroll
  for (int p = 0; p < 4; ++p) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int i = i0 + ty * 4 + p, j = j0 + tx * 4 + q;
      if (i < D && j < D) {
        atomicAdd(&sxx[(size_t)i * D + j], (double)acc[p][q]);
        if (ti != tj) atomicAdd(&sxx[(size_t)j * D + i], (double)acc[p][q]);
      }
    }
  }
  if (ti == tj && threadIdx.x < kFidTile && i0 + (int)threadIdx.x < D) atomicAdd(&sum[i0 + threadIdx.x], (double)colsum);
}

}  // namespace dm

using namespace dm;

extern "C" int dm_fk_reward(const float* qpos_dev, const float* ref_dev, int N, int H, float dt, float* body_pos_dev,
                            float* reward_dev, double* sums_dev, dm_stream_t st) {
  if (!qpos_dev || !ref_dev || !reward_dev || !sums_dev) return fail(DM_ERR_ARG, "null argument");
  if (N < 1 || H < 1 || !(dt > 0.f)) return fail(DM_ERR_ARG, "need N >= 1, H >= 1, dt > 0");
  int rc = check_device();
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  float* feat = nullptr;
  cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&feat), (size_t)H * kFeatFloats * sizeof(float), s);
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "cudaMallocAsync: %s", cudaGetErrorString(e));
  ref_features_kernel<<<(H + 63) / 64, 64, 0, s>>>(ref_dev, H, 1.f / dt, feat);
  const long long nfr = (long long)N * H;
  fk_reward_kernel<<<(unsigned)((nfr + kFkThreads - 1) / kFkThreads), kFkThreads, 0, s>>>(qpos_dev, feat, nfr, H, 1.f / dt,
                                                                                        body_pos_dev, reward_dev, sums_dev);
  e = cudaGetLastError();
  cudaFreeAsync(feat, s);
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "fk_reward launch: %s", cudaGetErrorString(e));
  return DM_OK;
}

extern "C" int dm_fid_moments(const float* X_dev, int N, int D, double* sum_dev, double* sxx_dev, dm_stream_t st) {
  if (!X_dev || !sum_dev || !sxx_dev) return fail(DM_ERR_ARG, "null argument");
  if (N < 1 || D < 1) return fail(DM_ERR_ARG, "need N >= 1, D >= 1");
  int rc = check_device();
  if (rc) return rc;
  const int nt = (D + kFidTile - 1) / kFidTile;
  const int tiles = nt * (nt + 1) / 2;
  int splits = (148 * 4 + tiles - 1) / tiles;
  if (splits < 1) splits = 1;
  int rows = (N + splits - 1) / splits;
  rows = (rows + kFidK - 1) / kFidK * kFidK;
  if (rows > 4096) rows = 4096;   // bound the fp32 accumulation length per partial sum
  splits = (N + rows - 1) / rows;
  dim3 grid(tiles, 1, splits);
  fid_moments_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(st)>>>(X_dev, N, D, rows, sum_dev, sxx_dev);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(DM_ERR_CUDA, "fid_moments launch: %s", cudaGetErrorString(e));
  return DM_OK;
}
