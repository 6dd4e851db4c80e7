// K5 arithmetic: forward kinematics of the DeepMimic humanoid + tracking-reward terms for ONE frame.
//
// FK follows MuJoCo mj_kinematics on /root/reference/diffusion/assets/dp_env_v2.xml:20-107 (the reference only
// reaches FK through mujoco_py sim.forward(), mocap_player.py:67-73): free root = qpos[0:7]; a child body starts at
// parent_pos + R_parent * body_offset with the parent's orientation; each hinge, in document order, rotates the
// frame about its body-frame axis through its anchor.  Reward: SURVEY.md 8(a) a15 (weights utils/mocap_util.py:26-29).
//
// Written __host__ __device__ so that tests/test_fk_host.py can run the very same arithmetic on the CPU against
// the float64 oracle (the shipped path is the CUDA kernel in score_kernels.cu; there is no CPU product path).
//
// Cost model (the kernel is issue-bound, not HBM-bound): 28 polynomial sin/cos pairs (17 FP32 ops each, accurate
// to 1 ulp so that FK stays inside the 1e-5 m bound), 12 quaternion products, 9 rotation angles through the
// cancellation-free identity 1 - |<f,g>| = |f -+ g|^2 / 2 and a polynomial for acos^2 (no atan2, no sqrt).
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define DM_HD __host__ __device__ __forceinline__
#else
#define DM_HD inline
#endif

namespace dm {
namespace fk {

constexpr int kFeatStride = 92;   // floats per reference frame (86 used; 23 x 16 B so that float4 rows do not conflict)
// feature layout (float offsets)
constexpr int F_ROOTQ = 0, F_JQ = 4, F_JA = 36, F_RATE = 40, F_EE = 68, F_COM = 80, F_ROOTW = 83;
constexpr int kBodies = 15;       // BODY_DEFS order (utils/mocap_util.py:17-19): 13 bodies + 2 wrist geoms
enum { B_ROOT, B_CHEST, B_NECK, B_RHIP, B_RKNEE, B_RANKLE, B_RSHO, B_RELB, B_RWRIST, B_LHIP, B_LKNEE, B_LANKLE, B_LSHO,
       B_LELB, B_LWRIST };

struct Q4 { float w, x, y, z; };
struct V3 { float x, y, z; };
struct M3 { V3 c1, c2, c3; };   // rotation matrix by columns

DM_HD float fast_rsqrt(float x) {
#if defined(__CUDA_ARCH__)
  return rsqrtf(x);
#else
  return 1.f / sqrtf(x);
#endif
}
DM_HD float fast_exp(float x) {
#if defined(__CUDA_ARCH__)
  return __expf(x);
#else
  return expf(x);
#endif
}

// sin(th/2), cos(th/2).  th is first reduced to [-pi, pi] (an odd number of turns flips the sign of both, i.e.
// gives -q, the same rotation); minimax polynomials on [-pi/2, pi/2]: |err| < 5e-9 (sin), 3e-10 (cos) before rounding.
DM_HD void sincos_half(float th, float& s, float& c) {
  const float k = rintf(th * 0.15915494309189535f);
  float t = fmaf(k, -6.2831854820251465f, th);
  t = fmaf(k, 1.7484555e-7f, t);
  const float r = 0.5f * t;
#if defined(__CUDA_ARCH__)
  // device: the SFU's sin / cos on the reduced half angle (|r| <= pi/2; absolute error ~4e-7 each): 4 instructions instead of
  // the 13 of the two polynomials.  FK positions stay within 3e-6 m of the float64 oracle (bar: 1e-5 m).
  s = __sinf(r);
  c = __cosf(r);
  return;
#endif
  const float z = r * r;
  float ps = 2.6052249021813675e-06f;
  ps = fmaf(ps, z, -0.00019809075284523637f);
  ps = fmaf(ps, z, 0.008333051063340868f);
  ps = fmaf(ps, z, -0.16666657991908637f);
  ps = ps * z;
  s = fmaf(ps, r, r);
  float pc = -2.6051495114016125e-07f;
  pc = fmaf(pc, z, 2.4760161346336438e-05f);
  pc = fmaf(pc, z, -0.0013888361400142136f);
  pc = fmaf(pc, z, 0.04166663625805913f);
  pc = fmaf(pc, z, -0.49999999358471525f);
  c = fmaf(pc, z, 1.0f);
}

DM_HD Q4 qmul(Q4 a, Q4 b) {
  Q4 r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  return r;
}
// conj(a) * b
DM_HD Q4 qmul_conj(Q4 a, Q4 b) {
  Q4 r;
  r.w = a.w * b.w + a.x * b.x + a.y * b.y + a.z * b.z;
  r.x = a.w * b.x - a.x * b.w - a.y * b.z + a.z * b.y;
  r.y = a.w * b.y + a.x * b.z - a.y * b.w - a.z * b.x;
  r.z = a.w * b.z - a.x * b.y + a.y * b.x - a.z * b.w;
  return r;
}
// a * (c, 0, -s, 0): hinge about (0,-1,0) (elbows, knees)
DM_HD Q4 qmul_negy(Q4 a, float s, float c) {
  Q4 r;
  r.w = a.w * c + a.y * s;
  r.x = a.x * c + a.z * s;
  r.y = a.y * c - a.w * s;
  r.z = a.z * c - a.x * s;
  return r;
}
DM_HD M3 qmat(Q4 q) {
  const float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
  const float xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z;
  const float wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
  M3 m;
  m.c1.x = fmaf(-2.f, yy + zz, 1.f); m.c1.y = 2.f * (xy + wz);          m.c1.z = 2.f * (xz - wy);
  m.c2.x = 2.f * (xy - wz);          m.c2.y = fmaf(-2.f, xx + zz, 1.f); m.c2.z = 2.f * (yz + wx);
  m.c3.x = 2.f * (xz + wy);          m.c3.y = 2.f * (yz - wx);          m.c3.z = fmaf(-2.f, xx + yy, 1.f);
  return m;
}
DM_HD V3 madd(V3 p, const M3& m, float vx, float vy, float vz) {   // p + M (vx, vy, vz)
  V3 r;
  r.x = fmaf(m.c1.x, vx, fmaf(m.c2.x, vy, fmaf(m.c3.x, vz, p.x)));
  r.y = fmaf(m.c1.y, vx, fmaf(m.c2.y, vy, fmaf(m.c3.y, vz, p.y)));
  r.z = fmaf(m.c1.z, vx, fmaf(m.c2.z, vy, fmaf(m.c3.z, vz, p.z)));
  return r;
}
DM_HD V3 madd_xz(V3 p, const M3& m, float vx, float vz) {
  V3 r;
  r.x = fmaf(m.c1.x, vx, fmaf(m.c3.x, vz, p.x));
  r.y = fmaf(m.c1.y, vx, fmaf(m.c3.y, vz, p.y));
  r.z = fmaf(m.c1.z, vx, fmaf(m.c3.z, vz, p.z));
  return r;
}
DM_HD V3 madd_z(V3 p, const M3& m, float vz) {
  V3 r;
  r.x = fmaf(m.c3.x, vz, p.x); r.y = fmaf(m.c3.y, vz, p.y); r.z = fmaf(m.c3.z, vz, p.z);
  return r;
}

// squared rotation angle between unit quaternions f and g: angle = 2 acos|<f,g>|.  u = 1 - |<f,g>| is computed as
// |f - sign g|^2 / 2 (no cancellation near u = 0); angle^2 = u P(u), P minimax on [0,1], relative error 5e-8.
DM_HD float angle_sq(Q4 f, float gw, float gx, float gy, float gz) {
  const float d = f.w * gw + f.x * gx + f.y * gy + f.z * gz;
  const float sg = d < 0.f ? -1.f : 1.f;
  const float dw = fmaf(-sg, gw, f.w), dx = fmaf(-sg, gx, f.x), dy = fmaf(-sg, gy, f.y), dz = fmaf(-sg, gz, f.z);
  float u = 0.5f * (dw * dw + dx * dx + dy * dy + dz * dz);
  u = fminf(u, 1.f);
  float p = 0.020827720299410272f;
  p = fmaf(p, u, -0.03622478550208995f);
  p = fmaf(p, u, 0.058430759845214404f);
  p = fmaf(p, u, 0.017858358457694793f);
  p = fmaf(p, u, 0.12063532375657782f);
  p = fmaf(p, u, 0.3547001425343587f);
  p = fmaf(p, u, 1.333376881657014f);
  p = fmaf(p, u, 7.999999637183414f);
  return p * u;
}

// local rotation of a 3-DoF joint (hinges x, y, z in document order): qx * qy * qz; also returns the half-angle
// sin / cos of the x and y hinges, needed when the y hinge is anchored off the body origin.
DM_HD Q4 euler_local(float ax, float ay, float az, float& sx, float& cx, float& sy, float& cy) {
  float sz, cz;
  sincos_half(ax, sx, cx);
  sincos_half(ay, sy, cy);
  sincos_half(az, sz, cz);
  const float w = cx * cy, x = sx * cy, y = cx * sy, z = sx * sy;
  Q4 r;
  r.w = w * cz - z * sz;
  r.x = x * cz + y * sz;
  r.y = y * cz - x * sz;
  r.z = z * cz + w * sz;
  return r;
}

DM_HD Q4 load_unit_quat(const float* q) {
  Q4 r; r.w = q[3]; r.x = q[4]; r.y = q[5]; r.z = q[6];
  const float n2 = r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z;
  if (n2 < 1e-20f) { r.w = 1.f; r.x = r.y = r.z = 0.f; }
  else { const float n = fast_rsqrt(n2); r.w *= n; r.x *= n; r.y *= n; r.z *= n; }
  return r;
}

// One frame.  q: the 35 qpos values of this frame, qp: of the previous frame (== q for the first frame of a clip,
// which gives zero velocity, mocap_v2.py:201-205).
//   kRef = true : write the frame's features to fo[kFeatStride] (the reference clip, once per call)
//   kRef = false: compare with the reference features g[] and return the four squared-error sums in acc[4]
//                 (pose, velocity, end effector, centre of mass)
//   bp          : optional world positions of the 15 bodies, 45 floats
template <bool kRef, bool kBody>
DM_HD void frame_pass(const float* q, const float* qp, float inv_dt, const float* g, float* fo, float* bp, float* acc) {
  float pose = 0.f, vel = 0.f, ee = 0.f;
  auto put_body = [&](int k, V3 p) {
    if (kBody) { bp[3 * k] = p.x; bp[3 * k + 1] = p.y; bp[3 * k + 2] = p.z; }
  };
  auto joint3 = [&](int k, float w, Q4 jq) {
    if (kRef) { fo[F_JQ + 4 * k] = jq.w; fo[F_JQ + 4 * k + 1] = jq.x; fo[F_JQ + 4 * k + 2] = jq.y; fo[F_JQ + 4 * k + 3] = jq.z; }
    else pose = fmaf(w, angle_sq(jq, g[F_JQ + 4 * k], g[F_JQ + 4 * k + 1], g[F_JQ + 4 * k + 2], g[F_JQ + 4 * k + 3]), pose);
  };
  auto joint1 = [&](int k, float w, float a) {
    if (kRef) fo[F_JA + k] = a;
    else { const float d = a - g[F_JA + k]; pose = fmaf(w * d, d, pose); }
  };
  auto effector = [&](int k, V3 p, V3 root) {
    const float dx = p.x - root.x, dy = p.y - root.y, dz = p.z - root.z;
    if (kRef) { fo[F_EE + 3 * k] = dx; fo[F_EE + 3 * k + 1] = dy; fo[F_EE + 3 * k + 2] = dz; }
    else {
      const float ex = dx - g[F_EE + 3 * k], ey = dy - g[F_EE + 3 * k + 1], ez = dz - g[F_EE + 3 * k + 2];
      ee += ex * ex + ey * ey + ez * ez;
    }
  };

  // ---- root ----
  const Q4 rq = load_unit_quat(q);
  const V3 rp = {q[0], q[1], q[2]};
  put_body(B_ROOT, rp);
  {
    const Q4 pq = load_unit_quat(qp);
    const Q4 d = qmul_conj(pq, rq);
    // axis * angle / dt (mocap_v2.py:155-178), shortest rotation
    const float vn2 = d.x * d.x + d.y * d.y + d.z * d.z;
    float wx = 0.f, wy = 0.f, wz = 0.f;
    if (vn2 >= 1e-24f) {
      const float vn = sqrtf(vn2);
      const float k = 2.f * atan2f(vn, fabsf(d.w)) * inv_dt * (d.w < 0.f ? -1.f : 1.f) / vn;
      wx = d.x * k; wy = d.y * k; wz = d.z * k;
    }
    if (kRef) {
      fo[F_ROOTQ] = rq.w; fo[F_ROOTQ + 1] = rq.x; fo[F_ROOTQ + 2] = rq.y; fo[F_ROOTQ + 3] = rq.z;
      fo[F_ROOTW] = wx; fo[F_ROOTW + 1] = wy; fo[F_ROOTW + 2] = wz;
    } else {
      pose = angle_sq(rq, g[F_ROOTQ], g[F_ROOTQ + 1], g[F_ROOTQ + 2], g[F_ROOTQ + 3]);
      const float ex = wx - g[F_ROOTW], ey = wy - g[F_ROOTW + 1], ez = wz - g[F_ROOTW + 2];
      vel = ex * ex + ey * ey + ez * ez;
    }
  }
  // ---- joint rates: finite differences of the 28 hinge angles; weight of the joint each angle belongs to ----
  {
    const float wr[28] = {0.5f, 0.5f, 0.5f, 0.3f, 0.3f, 0.3f, 0.3f, 0.3f, 0.3f, 0.2f, 0.3f, 0.3f, 0.3f, 0.2f,
                          0.5f, 0.5f, 0.5f, 0.3f, 0.2f, 0.2f, 0.2f, 0.5f, 0.5f, 0.5f, 0.3f, 0.2f, 0.2f, 0.2f};
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int d = 0; d < 28; ++d) {
      const float r = (q[7 + d] - qp[7 + d]) * inv_dt;
      if (kRef) fo[F_RATE + d] = r;
      else { const float e = r - g[F_RATE + d]; vel = fmaf(wr[d] * e, e, vel); }
    }
  }
  const M3 rm = qmat(rq);
  // centre of mass: sum over bodies of m p + R (m c), geoms of dp_env_v2.xml (capsule centre = midpoint of fromto)
  V3 com = {2.5f * rp.x, 2.5f * rp.y, 2.5f * rp.z};
  com = madd_z(com, rm, 2.5f * 0.07f);

  float sx, cx, sy, cy;
  // ---- chest (qpos 7-9), neck (10-12) ----
  const V3 chest_p = madd_z(rp, rm, 0.236151f);
  Q4 jq = euler_local(q[7], q[8], q[9], sx, cx, sy, cy);
  joint3(0, 0.5f, jq);
  const Q4 chest_q = qmul(rq, jq);
  const M3 cm = qmat(chest_q);
  put_body(B_CHEST, chest_p);
  com.x = fmaf(3.f, chest_p.x, com.x); com.y = fmaf(3.f, chest_p.y, com.y); com.z = fmaf(3.f, chest_p.z, com.z);
  com = madd_z(com, cm, 3.f * 0.12f);
  {
    const V3 neck_p = madd_z(chest_p, cm, 0.223894f);
    jq = euler_local(q[10], q[11], q[12], sx, cx, sy, cy);
    joint3(1, 0.3f, jq);
    const M3 nm = qmat(qmul(chest_q, jq));
    put_body(B_NECK, neck_p);
    com.x = fmaf(2.f, neck_p.x, com.x); com.y = fmaf(2.f, neck_p.y, com.y); com.z = fmaf(2.f, neck_p.z, com.z);
    com = madd_z(com, nm, 2.f * 0.175f);
  }
  // ---- arms: shoulder (13-15 / 17-19), elbow (16 / 20): axis (0,-1,0), anchor (0.028,0,0); wrist geom at (0,0,-0.258947)
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int s = 0; s < 2; ++s) {
    const float ysign = s == 0 ? -1.f : 1.f;
    const int qi = s == 0 ? 13 : 17;
    const V3 sh_p = madd(chest_p, cm, -0.02405f, ysign * 0.18311f, 0.2435f);
    jq = euler_local(q[qi], q[qi + 1], q[qi + 2], sx, cx, sy, cy);
    joint3(2 + s, 0.3f, jq);
    const Q4 sh_q = qmul(chest_q, jq);
    const M3 sm = qmat(sh_q);
    put_body(s == 0 ? B_RSHO : B_LSHO, sh_p);
    const float ea = q[qi + 3];
    joint1(s, 0.2f, ea);
    float se, ce;
    sincos_half(ea, se, ce);
    // elbow origin: offset (0,0,-0.274788), then the anchored hinge moves it by R (anc - R_hinge anc), anc = (0.028,0,0)
    const V3 el_p = madd_xz(sh_p, sm, 2.f * 0.028f * se * se, -0.274788f - 2.f * 0.028f * se * ce);
    const M3 em = qmat(qmul_negy(sh_q, se, ce));
    put_body(s == 0 ? B_RELB : B_LELB, el_p);
    const V3 wr_p = madd_z(el_p, em, -0.258947f);
    put_body(s == 0 ? B_RWRIST : B_LWRIST, wr_p);
    effector(2 + s, wr_p, rp);
    com.x = fmaf(1.5f, sh_p.x, com.x); com.y = fmaf(1.5f, sh_p.y, com.y); com.z = fmaf(1.5f, sh_p.z, com.z);
    com = madd_z(com, sm, 1.5f * -0.14f);
    com.x = fmaf(1.5f, el_p.x, com.x); com.y = fmaf(1.5f, el_p.y, com.y); com.z = fmaf(1.5f, el_p.z, com.z);
    com = madd_z(com, em, 1.0f * -0.12f + 0.5f * -0.258947f);
  }
  // ---- legs: hip (21-23 / 28-30; y hinge anchored at (0.02,0,0)), knee (24 / 31; axis (0,-1,0), anchor (-0.035,0,0)),
  //      ankle (25-27 / 32-34; y hinge anchored at (-0.02,0,0))
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int s = 0; s < 2; ++s) {
    const float ysign = s == 0 ? -1.f : 1.f;
    const int qi = s == 0 ? 21 : 28;
    jq = euler_local(q[qi], q[qi + 1], q[qi + 2], sx, cx, sy, cy);
    joint3(4 + 2 * s, 0.5f, jq);
    // anchored y hinge after the x hinge: shift R_parent R_x (dx, 0, dz), (dx, dz) = a (1 - cos, sin) of the y angle
    float dx = 2.f * 0.02f * sy * sy, dz = 2.f * 0.02f * sy * cy;
    float sinx = 2.f * sx * cx, cosx = fmaf(-2.f * sx, sx, 1.f);
    const V3 hip_p = madd(rp, rm, dx, fmaf(-dz, sinx, ysign * 0.094887f), fmaf(dz, cosx, -0.05f));
    const Q4 hip_q = qmul(rq, jq);
    const M3 hm = qmat(hip_q);
    put_body(s == 0 ? B_RHIP : B_LHIP, hip_p);
    const float ka = q[qi + 3];
    joint1(2 + s, 0.3f, ka);
    float sk, ck;
    sincos_half(ka, sk, ck);
    const V3 kn_p = madd_xz(hip_p, hm, 2.f * -0.035f * sk * sk, -0.421546f - 2.f * -0.035f * sk * ck);
    const Q4 kn_q = qmul_negy(hip_q, sk, ck);
    const M3 km = qmat(kn_q);
    put_body(s == 0 ? B_RKNEE : B_LKNEE, kn_p);
    jq = euler_local(q[qi + 4], q[qi + 5], q[qi + 6], sx, cx, sy, cy);
    joint3(5 + 2 * s, 0.2f, jq);
    dx = 2.f * -0.02f * sy * sy; dz = 2.f * -0.02f * sy * cy;
    sinx = 2.f * sx * cx; cosx = fmaf(-2.f * sx, sx, 1.f);
    const V3 an_p = madd(kn_p, km, dx, -dz * sinx, fmaf(dz, cosx, -0.40987f));
    const M3 am = qmat(qmul(kn_q, jq));
    put_body(s == 0 ? B_RANKLE : B_LANKLE, an_p);
    effector(s, an_p, rp);
    com.x = fmaf(4.5f, hip_p.x, com.x); com.y = fmaf(4.5f, hip_p.y, com.y); com.z = fmaf(4.5f, hip_p.z, com.z);
    com = madd_z(com, hm, 4.5f * -0.21f);
    com.x = fmaf(3.f, kn_p.x, com.x); com.y = fmaf(3.f, kn_p.y, com.y); com.z = fmaf(3.f, kn_p.z, com.z);
    com = madd_z(com, km, 3.f * -0.2f);
    com.x += an_p.x; com.y += an_p.y; com.z += an_p.z;
    com = madd_xz(com, am, 0.045f, -0.0425f);
  }
  const float inv_m = 1.f / 30.5f;
  com.x *= inv_m; com.y *= inv_m; com.z *= inv_m;
  if (kRef) {
    fo[F_COM] = com.x; fo[F_COM + 1] = com.y; fo[F_COM + 2] = com.z;
  } else {
    const float ex = com.x - g[F_COM], ey = com.y - g[F_COM + 1], ez = com.z - g[F_COM + 2];
    acc[0] = pose; acc[1] = vel; acc[2] = ee; acc[3] = ex * ex + ey * ey + ez * ez;
  }
}

}  // namespace fk
}  // namespace dm
