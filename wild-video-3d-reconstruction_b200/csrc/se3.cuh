// Minimal fp32 SE3 algebra for the fastba kernels (poses are [tx ty tz qx qy qz qw], world->camera).
// Same maps as the helpers in dpvo/fastba/ba_cuda.cu:18-156, written from the group definitions.
#pragma once
#include <cuda_runtime.h>

namespace wv {

struct Quat { float x, y, z, w; };
struct Vec3 { float x, y, z; };

__device__ __forceinline__ Vec3 cross(const Vec3& a, const Vec3& b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// R(q) v = v + w*(2 q x v) + q x (2 q x v)
__device__ __forceinline__ Vec3 rotate(const Quat& q, const Vec3& v) {
  const Vec3 qv{q.x, q.y, q.z};
  Vec3 u = cross(qv, v);
  u.x *= 2.f; u.y *= 2.f; u.z *= 2.f;
  const Vec3 c = cross(qv, u);
  return {v.x + q.w * u.x + c.x, v.y + q.w * u.y + c.y, v.z + q.w * u.z + c.z};
}

__device__ __forceinline__ Quat conj(const Quat& q) { return {-q.x, -q.y, -q.z, q.w}; }

// Hamilton product a (x) b
__device__ __forceinline__ Quat qmul(const Quat& a, const Quat& b) {
  return {a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
          a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
          a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x,
          a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z};
}

// G_ij = G_j o G_i^-1  (maps camera-i coordinates to camera-j coordinates)
__device__ __forceinline__ void relative_pose(const Vec3& ti, const Quat& qi, const Vec3& tj, const Quat& qj,
                                              Vec3& tij, Quat& qij) {
  qij = qmul(qj, conj(qi));
  const Vec3 r = rotate(qij, ti);
  tij = {tj.x - r.x, tj.y - r.y, tj.z - r.z};
}

// J_i = Ad(G_ij)^T-style pull-back of a 6-vector of pose derivatives (tau part, phi part) to frame i
__device__ __forceinline__ void adjoint_T(const Vec3& t, const Quat& q, const float* X, float* Y) {
  const Quat qi = conj(q);
  const Vec3 a{X[0], X[1], X[2]}, b{X[3], X[4], X[5]};
  const Vec3 ra = rotate(qi, a);
  const Vec3 rb = rotate(qi, b);
  const Vec3 u = cross(a, t);  // = -(t x a)
  const Vec3 ru = rotate(qi, u);
  Y[0] = ra.x; Y[1] = ra.y; Y[2] = ra.z;
  Y[3] = rb.x + ru.x; Y[4] = rb.y + ru.y; Y[5] = rb.z + ru.z;
}

// SO3 exponential as a unit quaternion; Taylor branch below theta^2 = 1e-8 (ba_cuda.cu:80-86)
__device__ __forceinline__ Quat so3_exp(const Vec3& phi) {
  const float th2 = phi.x * phi.x + phi.y * phi.y + phi.z * phi.z;
  float imag, real;
  if (th2 < 1e-8f) {
    const float th4 = th2 * th2;
    imag = 0.5f - (1.f / 48.f) * th2 + (1.f / 3840.f) * th4;
    real = 1.f - (1.f / 8.f) * th2 + (1.f / 384.f) * th4;
  } else {
    const float th = sqrtf(th2);
    imag = sinf(0.5f * th) / th;
    real = cosf(0.5f * th);
  }
  return {imag * phi.x, imag * phi.y, imag * phi.z, real};
}

// T <- Exp(xi) o T, xi = (tau, phi); V-matrix terms only above theta = 1e-4 (ba_cuda.cu:124-136)
__device__ __forceinline__ void se3_retract(const float* xi, Vec3& t, Quat& q) {
  const Vec3 tau{xi[0], xi[1], xi[2]}, phi{xi[3], xi[4], xi[5]};
  const Quat dq = so3_exp(phi);
  Vec3 dt = tau;
  const float th2 = phi.x * phi.x + phi.y * phi.y + phi.z * phi.z;
  const float th = sqrtf(th2);
  if (th > 1e-4f) {
    const float a = (1.f - cosf(th)) / th2;
    const float b = (th - sinf(th)) / (th * th2);
    const Vec3 c1 = cross(phi, tau);
    const Vec3 c2 = cross(phi, c1);
    dt.x += a * c1.x + b * c2.x;
    dt.y += a * c1.y + b * c2.y;
    dt.z += a * c1.z + b * c2.z;
  }
  const Vec3 rt = rotate(dq, t);
  t = {rt.x + dt.x, rt.y + dt.y, rt.z + dt.z};
  q = qmul(dq, q);
}

}  // namespace wv
