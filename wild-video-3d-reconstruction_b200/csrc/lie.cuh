// Lie-group algebra behind `lietorch_backends` (SO3, RxSO3, SE3, Sim3): plain fixed-size arrays, host + device, no Eigen.
//
// The contract is the reference's group arithmetic, dpvo/lietorch/include/{so3,rxso3,se3,sim3}.h -- storage layouts
// (quaternion as x, y, z, w; translation first; scale last), re-normalisation of the quaternion on every construction
// (so3.h:31-37), the Taylor branches below EPS = 1e-6 (common.h:7), the atan-based logarithm (so3.h:115-152), and the series
// forms of the Sim3 Jacobians including the term sim3.h:196 drops.  Everything here is written from those formulas; trigonometric
// coefficients are evaluated in double for both scalar types (the reference's float instantiation mixes float functions with
// double literals; double is at least as accurate).
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define LIE_HD __host__ __device__ __forceinline__
#else
#define LIE_HD inline
#endif

namespace lie {

constexpr double kEps = 1e-6;
constexpr double kPi = 3.14159265358979323846;

template <typename T, int R, int C>
struct Mat {
  T a[R * C];
  LIE_HD T& operator()(int r, int c) { return a[r * C + c]; }
  LIE_HD const T& operator()(int r, int c) const { return a[r * C + c]; }
  LIE_HD static Mat zero() {
    Mat m;
#pragma unroll
    for (int i = 0; i < R * C; ++i) m.a[i] = T(0);
    return m;
  }
  LIE_HD static Mat identity() {
    Mat m = zero();
#pragma unroll
    for (int i = 0; i < (R < C ? R : C); ++i) m(i, i) = T(1);
    return m;
  }
};

template <typename T, int N>
struct Vec {
  T a[N];
  LIE_HD T& operator[](int i) { return a[i]; }
  LIE_HD const T& operator[](int i) const { return a[i]; }
  LIE_HD static Vec zero() {
    Vec v;
#pragma unroll
    for (int i = 0; i < N; ++i) v.a[i] = T(0);
    return v;
  }
};

template <typename T, int R, int K, int C>
LIE_HD Mat<T, R, C> operator*(const Mat<T, R, K>& A, const Mat<T, K, C>& B) {
  Mat<T, R, C> o;
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int c = 0; c < C; ++c) {
      T s = T(0);
#pragma unroll
      for (int k = 0; k < K; ++k) s += A(r, k) * B(k, c);
      o(r, c) = s;
    }
  return o;
}
template <typename T, int R, int C>
LIE_HD Vec<T, R> operator*(const Mat<T, R, C>& A, const Vec<T, C>& v) {
  Vec<T, R> o;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    T s = T(0);
#pragma unroll
    for (int c = 0; c < C; ++c) s += A(r, c) * v[c];
    o[r] = s;
  }
  return o;
}
// row vector times matrix
template <typename T, int R, int C>
LIE_HD Vec<T, C> rowmul(const Vec<T, R>& v, const Mat<T, R, C>& A) {
  Vec<T, C> o;
#pragma unroll
  for (int c = 0; c < C; ++c) {
    T s = T(0);
#pragma unroll
    for (int r = 0; r < R; ++r) s += v[r] * A(r, c);
    o[c] = s;
  }
  return o;
}
template <typename T, int R, int C>
LIE_HD Mat<T, R, C> operator+(const Mat<T, R, C>& A, const Mat<T, R, C>& B) {
  Mat<T, R, C> o;
#pragma unroll
  for (int i = 0; i < R * C; ++i) o.a[i] = A.a[i] + B.a[i];
  return o;
}
template <typename T, int R, int C>
LIE_HD Mat<T, R, C> operator-(const Mat<T, R, C>& A, const Mat<T, R, C>& B) {
  Mat<T, R, C> o;
#pragma unroll
  for (int i = 0; i < R * C; ++i) o.a[i] = A.a[i] - B.a[i];
  return o;
}
template <typename T, int R, int C>
LIE_HD Mat<T, R, C> operator*(T s, const Mat<T, R, C>& A) {
  Mat<T, R, C> o;
#pragma unroll
  for (int i = 0; i < R * C; ++i) o.a[i] = s * A.a[i];
  return o;
}
template <typename T, int N>
LIE_HD Vec<T, N> operator*(T s, const Vec<T, N>& v) {
  Vec<T, N> o;
#pragma unroll
  for (int i = 0; i < N; ++i) o[i] = s * v[i];
  return o;
}
template <typename T, int N>
LIE_HD Vec<T, N> operator+(const Vec<T, N>& a, const Vec<T, N>& b) {
  Vec<T, N> o;
#pragma unroll
  for (int i = 0; i < N; ++i) o[i] = a[i] + b[i];
  return o;
}
template <typename T, int N>
LIE_HD Vec<T, N> operator-(const Vec<T, N>& a) {
  Vec<T, N> o;
#pragma unroll
  for (int i = 0; i < N; ++i) o[i] = -a[i];
  return o;
}
template <typename T, int R, int C>
LIE_HD Mat<T, C, R> transpose(const Mat<T, R, C>& A) {
  Mat<T, C, R> o;
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int c = 0; c < C; ++c) o(c, r) = A(r, c);
  return o;
}
// write block B into A at (r0, c0)
template <typename T, int R, int C, int BR, int BC>
LIE_HD void put(Mat<T, R, C>& A, int r0, int c0, const Mat<T, BR, BC>& B) {
#pragma unroll
  for (int r = 0; r < BR; ++r)
#pragma unroll
    for (int c = 0; c < BC; ++c) A(r0 + r, c0 + c) = B(r, c);
}
template <typename T, int R, int C>
LIE_HD void put_col(Mat<T, R, C>& A, int r0, int c, const Vec<T, 3>& v) {
  A(r0, c) = v[0]; A(r0 + 1, c) = v[1]; A(r0 + 2, c) = v[2];
}

template <typename T> using V3 = Vec<T, 3>;
template <typename T> using M3 = Mat<T, 3, 3>;

template <typename T>
LIE_HD V3<T> cross(const V3<T>& a, const V3<T>& b) {
  V3<T> o;
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
  return o;
}
template <typename T>
LIE_HD M3<T> hat(const V3<T>& p) {
  M3<T> m;
  m(0, 0) = T(0); m(0, 1) = -p[2]; m(0, 2) = p[1];
  m(1, 0) = p[2]; m(1, 1) = T(0); m(1, 2) = -p[0];
  m(2, 0) = -p[1]; m(2, 1) = p[0]; m(2, 2) = T(0);
  return m;
}
template <typename T>
LIE_HD M3<T> inverse3(const M3<T>& m) {
  const T c00 = m(1, 1) * m(2, 2) - m(1, 2) * m(2, 1);
  const T c01 = m(1, 2) * m(2, 0) - m(1, 0) * m(2, 2);
  const T c02 = m(1, 0) * m(2, 1) - m(1, 1) * m(2, 0);
  const T inv = T(1) / (m(0, 0) * c00 + m(0, 1) * c01 + m(0, 2) * c02);
  M3<T> o;
  o(0, 0) = c00 * inv;
  o(0, 1) = (m(0, 2) * m(2, 1) - m(0, 1) * m(2, 2)) * inv;
  o(0, 2) = (m(0, 1) * m(1, 2) - m(0, 2) * m(1, 1)) * inv;
  o(1, 0) = c01 * inv;
  o(1, 1) = (m(0, 0) * m(2, 2) - m(0, 2) * m(2, 0)) * inv;
  o(1, 2) = (m(0, 2) * m(1, 0) - m(0, 0) * m(1, 2)) * inv;
  o(2, 0) = c02 * inv;
  o(2, 1) = (m(0, 1) * m(2, 0) - m(0, 0) * m(2, 1)) * inv;
  o(2, 2) = (m(0, 0) * m(1, 1) - m(0, 1) * m(1, 0)) * inv;
  return o;
}

// unit quaternion, stored (x, y, z, w)
template <typename T>
struct Quat {
  T x, y, z, w;
  LIE_HD V3<T> vec() const { V3<T> v; v[0] = x; v[1] = y; v[2] = z; return v; }
  LIE_HD void normalize() {
    const T n = sqrt(x * x + y * y + z * z + w * w);
    x /= n; y /= n; z /= n; w /= n;
  }
  LIE_HD Quat conj() const { return Quat{-x, -y, -z, w}; }
  LIE_HD Quat operator*(const Quat& b) const {
    return Quat{w * b.x + x * b.w + y * b.z - z * b.y, w * b.y + y * b.w + z * b.x - x * b.z,
                w * b.z + z * b.w + x * b.y - y * b.x, w * b.w - x * b.x - y * b.y - z * b.z};
  }
  LIE_HD V3<T> rotate(const V3<T>& p) const {
    V3<T> uv = cross(vec(), p);
    uv = uv + uv;
    return p + w * uv + cross(vec(), uv);
  }
  LIE_HD M3<T> matrix() const {
    const T tx = T(2) * x, ty = T(2) * y, tz = T(2) * z;
    const T twx = tx * w, twy = ty * w, twz = tz * w;
    const T txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
    M3<T> R;
    R(0, 0) = T(1) - (tyy + tzz); R(0, 1) = txy - twz; R(0, 2) = txz + twy;
    R(1, 0) = txy + twz; R(1, 1) = T(1) - (txx + tzz); R(1, 2) = tyz - twx;
    R(2, 0) = txz - twy; R(2, 1) = tyz + twx; R(2, 2) = T(1) - (txx + tyy);
    return R;
  }
};

// rotation-vector helpers shared by the groups --------------------------------------------------------------------------------
template <typename T>
LIE_HD Quat<T> quat_exp(const V3<T>& phi) {
  const double t2 = (double)phi[0] * phi[0] + (double)phi[1] * phi[1] + (double)phi[2] * phi[2];
  const double th = sqrt(t2);
  double im, re;
  if (th < kEps) {
    const double t4 = t2 * t2;
    im = 0.5 - (1.0 / 48.0) * t2 + (1.0 / 3840.0) * t4;
    re = 1.0 - (1.0 / 8.0) * t2 + (1.0 / 384.0) * t4;
  } else {
    im = sin(0.5 * th) / th;
    re = cos(0.5 * th);
  }
  Quat<T> q{T(im * phi[0]), T(im * phi[1]), T(im * phi[2]), T(re)};
  q.normalize();
  return q;
}
template <typename T>
LIE_HD V3<T> quat_log(const Quat<T>& q) {
  const double n2 = (double)q.x * q.x + (double)q.y * q.y + (double)q.z * q.z;
  const double w = q.w;
  double f;
  if (n2 < kEps * kEps) {
    f = 2.0 / w - (2.0 / 3.0) * n2 / (w * w * w);
  } else {
    const double n = sqrt(n2);
    if (fabs(w) < kEps) f = (w > 0.0 ? kPi : -kPi) / n;
    else f = 2.0 * atan(n / w) / n;
  }
  V3<T> o;
  o[0] = T(f * q.x); o[1] = T(f * q.y); o[2] = T(f * q.z);
  return o;
}
template <typename T>
LIE_HD M3<T> so3_left_jacobian(const V3<T>& phi) {
  const M3<T> Phi = hat(phi);
  const M3<T> Phi2 = Phi * Phi;
  const double t2 = (double)phi[0] * phi[0] + (double)phi[1] * phi[1] + (double)phi[2] * phi[2];
  const double th = sqrt(t2);
  const double c1 = th < kEps ? 0.5 - (1.0 / 24.0) * t2 : (1.0 - cos(th)) / t2;
  const double c2 = th < kEps ? 1.0 / 6.0 - (1.0 / 120.0) * t2 : (th - sin(th)) / (t2 * th);
  return M3<T>::identity() + T(c1) * Phi + T(c2) * Phi2;
}
template <typename T>
LIE_HD M3<T> so3_left_jacobian_inverse(const V3<T>& phi) {
  const M3<T> Phi = hat(phi);
  const M3<T> Phi2 = Phi * Phi;
  const double t2 = (double)phi[0] * phi[0] + (double)phi[1] * phi[1] + (double)phi[2] * phi[2];
  const double th = sqrt(t2);
  const double h = 0.5 * th;
  const double c2 = th < kEps ? 1.0 / 12.0 : (1.0 - th * cos(h) / (2.0 * sin(h))) / (th * th);
  return M3<T>::identity() + T(-0.5) * Phi + T(c2) * Phi2;
}
// 4 x 4 projector block of a unit quaternion (so3.h:82-93): rows 0..2 = (w I + hat(-v)) / 2, row 3 = -v / 2, last column zero
template <typename T, int N>
LIE_HD void put_quat_projector(Mat<T, N, N>& J, int o, const Quat<T>& q) {
  const M3<T> H = hat(-q.vec());
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) J(o + r, o + c) = T(0.5) * ((r == c ? q.w : T(0)) + H(r, c));
  J(o + 3, o + 0) = T(0.5) * -q.x; J(o + 3, o + 1) = T(0.5) * -q.y; J(o + 3, o + 2) = T(0.5) * -q.z;
}

// ------------------------------------------------------------------------------------------------------------------ SO3 ----
template <typename T>
struct SO3 {
  static constexpr int K = 3, N = 4;
  using Tangent = Vec<T, K>;
  using Adjoint = Mat<T, K, K>;
  Quat<T> q;
  LIE_HD SO3() : q{T(0), T(0), T(0), T(1)} {}
  LIE_HD explicit SO3(const Quat<T>& q_) : q(q_) { q.normalize(); }
  LIE_HD explicit SO3(const T* d) : q{d[0], d[1], d[2], d[3]} { q.normalize(); }
  LIE_HD void store(T* d) const { d[0] = q.x; d[1] = q.y; d[2] = q.z; d[3] = q.w; }
  LIE_HD SO3 inv() const { return SO3(q.conj()); }
  LIE_HD SO3 operator*(const SO3& o) const { return SO3(q * o.q); }
  LIE_HD V3<T> act(const V3<T>& p) const { return q.rotate(p); }
  LIE_HD Vec<T, 4> act4(const Vec<T, 4>& p) const {
    V3<T> p3; p3[0] = p[0]; p3[1] = p[1]; p3[2] = p[2];
    const V3<T> r = act(p3);
    Vec<T, 4> o; o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = p[3];
    return o;
  }
  LIE_HD Adjoint Adj() const { return q.matrix(); }
  LIE_HD Mat<T, 4, 4> matrix4() const { Mat<T, 4, 4> M = Mat<T, 4, 4>::identity(); put(M, 0, 0, q.matrix()); return M; }
  LIE_HD Mat<T, N, N> projector() const { Mat<T, N, N> J = Mat<T, N, N>::zero(); put_quat_projector(J, 0, q); return J; }
  LIE_HD static Adjoint adj(const Tangent& a) { return hat(a); }
  LIE_HD Tangent Log() const { return quat_log(q); }
  LIE_HD static SO3 Exp(const Tangent& a) { SO3 o; o.q = quat_exp(a); return o; }
  LIE_HD static Adjoint left_jacobian(const Tangent& a) { return so3_left_jacobian(a); }
  LIE_HD static Adjoint left_jacobian_inverse(const Tangent& a) { return so3_left_jacobian_inverse(a); }
  LIE_HD static Mat<T, 3, K> act_jacobian(const V3<T>& p) { return hat(-p); }
  LIE_HD static Mat<T, 4, K> act4_jacobian(const Vec<T, 4>& p) {
    Mat<T, 4, K> J = Mat<T, 4, K>::zero();
    V3<T> p3; p3[0] = p[0]; p3[1] = p[1]; p3[2] = p[2];
    put(J, 0, 0, hat(-p3));
    return J;
  }
};

// ---------------------------------------------------------------------------------------------------------------- RxSO3 ----
template <typename T>
struct RxSO3 {
  static constexpr int K = 4, N = 5;
  using Tangent = Vec<T, K>;
  using Adjoint = Mat<T, K, K>;
  Quat<T> q;
  T s;
  LIE_HD RxSO3() : q{T(0), T(0), T(0), T(1)}, s(T(1)) {}
  LIE_HD RxSO3(const Quat<T>& q_, T s_) : q(q_), s(s_) { q.normalize(); }
  LIE_HD explicit RxSO3(const T* d) : q{d[0], d[1], d[2], d[3]}, s(d[4]) { q.normalize(); }
  LIE_HD void store(T* d) const { d[0] = q.x; d[1] = q.y; d[2] = q.z; d[3] = q.w; d[4] = s; }
  LIE_HD RxSO3 inv() const { return RxSO3(q.conj(), T(1.0 / s)); }
  LIE_HD RxSO3 operator*(const RxSO3& o) const { return RxSO3(q * o.q, s * o.s); }
  LIE_HD V3<T> act(const V3<T>& p) const { return s * q.rotate(p); }
  LIE_HD Vec<T, 4> act4(const Vec<T, 4>& p) const {
    V3<T> p3; p3[0] = p[0]; p3[1] = p[1]; p3[2] = p[2];
    const V3<T> r = act(p3);
    Vec<T, 4> o; o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = p[3];
    return o;
  }
  LIE_HD M3<T> rotation() const { return q.matrix(); }
  LIE_HD M3<T> matrix3() const { return s * q.matrix(); }
  LIE_HD Adjoint Adj() const { Adjoint A = Adjoint::identity(); put(A, 0, 0, q.matrix()); return A; }
  LIE_HD Mat<T, 4, 4> matrix4() const { Mat<T, 4, 4> M = Mat<T, 4, 4>::identity(); put(M, 0, 0, matrix3()); return M; }
  LIE_HD Mat<T, N, N> projector() const {
    Mat<T, N, N> J = Mat<T, N, N>::zero();
    put_quat_projector(J, 0, q);
    J(4, 3) = s;
    return J;
  }
  LIE_HD static Adjoint adj(const Tangent& a) {
    Adjoint A = Adjoint::zero();
    V3<T> phi; phi[0] = a[0]; phi[1] = a[1]; phi[2] = a[2];
    put(A, 0, 0, hat(phi));
    return A;
  }
  LIE_HD Tangent Log() const {
    const V3<T> phi = quat_log(q);
    Tangent o; o[0] = phi[0]; o[1] = phi[1]; o[2] = phi[2]; o[3] = T(log((double)s));
    return o;
  }
  LIE_HD static RxSO3 Exp(const Tangent& a) {
    V3<T> phi; phi[0] = a[0]; phi[1] = a[1]; phi[2] = a[2];
    RxSO3 o; o.q = quat_exp(phi); o.s = T(exp((double)a[3]));
    return o;
  }
  // W of the Sim3 exponential (rxso3.h:188-229)
  LIE_HD static M3<T> calcW(const Tangent& a) {
    V3<T> phi; phi[0] = a[0]; phi[1] = a[1]; phi[2] = a[2];
    const double sigma = a[3];
    const double th = sqrt((double)phi[0] * phi[0] + (double)phi[1] * phi[1] + (double)phi[2] * phi[2]);
    const M3<T> Phi = hat(phi);
    const M3<T> Phi2 = Phi * Phi;
    const double scale = exp(sigma);
    double A, B, C;
    if (fabs(sigma) < kEps) {
      C = 1.0;
      if (fabs(th) < kEps) { A = 0.5; B = 1.0 / 6.0; }
      else { const double t2 = th * th; A = (1.0 - cos(th)) / t2; B = (th - sin(th)) / (t2 * th); }
    } else {
      C = (scale - 1.0) / sigma;
      if (fabs(th) < kEps) {
        const double s2 = sigma * sigma;
        A = ((sigma - 1.0) * scale + 1.0) / s2;
        B = (scale * 0.5 * s2 + scale - 1.0 - sigma * scale) / (s2 * sigma);
      } else {
        const double t2 = th * th;
        const double a_ = scale * sin(th), b_ = scale * cos(th), c_ = t2 + sigma * sigma;
        A = (a_ * sigma + (1.0 - b_) * th) / (th * c_);
        B = (C - ((b_ - 1.0) * sigma + a_ * th) / c_) * 1.0 / t2;
      }
    }
    return T(A) * Phi + T(B) * Phi2 + T(C) * M3<T>::identity();
  }
  LIE_HD static Adjoint left_jacobian(const Tangent& a) {
    Adjoint J = Adjoint::identity();
    V3<T> phi; phi[0] = a[0]; phi[1] = a[1]; phi[2] = a[2];
    put(J, 0, 0, so3_left_jacobian(phi));
    return J;
  }
  LIE_HD static Adjoint left_jacobian_inverse(const Tangent& a) {
    Adjoint J = Adjoint::identity();
    V3<T> phi; phi[0] = a[0]; phi[1] = a[1]; phi[2] = a[2];
    put(J, 0, 0, so3_left_jacobian_inverse(phi));
    return J;
  }
  LIE_HD static Mat<T, 3, K> act_jacobian(const V3<T>& p) {
    Mat<T, 3, K> J;
    put(J, 0, 0, hat(-p));
    put_col(J, 0, 3, p);
    return J;
  }
  LIE_HD static Mat<T, 4, K> act4_jacobian(const Vec<T, 4>& p) {
    Mat<T, 4, K> J = Mat<T, 4, K>::zero();
    V3<T> p3; p3[0] = p[0]; p3[1] = p[1]; p3[2] = p[2];
    put(J, 0, 0, hat(-p3));
    put_col(J, 0, 3, p3);
    return J;
  }
};

// ------------------------------------------------------------------------------------------------------------------ SE3 ----
template <typename T>
struct SE3 {
  static constexpr int K = 6, N = 7;
  using Tangent = Vec<T, K>;
  using Adjoint = Mat<T, K, K>;
  V3<T> t;
  SO3<T> r;
  LIE_HD SE3() : t(V3<T>::zero()) {}
  LIE_HD SE3(const SO3<T>& r_, const V3<T>& t_) : t(t_), r(r_) {}
  LIE_HD explicit SE3(const T* d) : r(d + 3) { t[0] = d[0]; t[1] = d[1]; t[2] = d[2]; }
  LIE_HD void store(T* d) const { d[0] = t[0]; d[1] = t[1]; d[2] = t[2]; r.store(d + 3); }
  LIE_HD SE3 inv() const { const SO3<T> ri = r.inv(); return SE3(ri, -ri.act(t)); }
  LIE_HD SE3 operator*(const SE3& o) const { return SE3(r * o.r, t + r.act(o.t)); }
  LIE_HD V3<T> act(const V3<T>& p) const { return r.act(p) + t; }
  LIE_HD Vec<T, 4> act4(const Vec<T, 4>& p) const {
    V3<T> p3; p3[0] = p[0]; p3[1] = p[1]; p3[2] = p[2];
    const V3<T> q = r.act(p3) + p[3] * t;
    Vec<T, 4> o; o[0] = q[0]; o[1] = q[1]; o[2] = q[2]; o[3] = p[3];
    return o;
  }
  LIE_HD Adjoint Adj() const {
    const M3<T> R = r.q.matrix();
    Adjoint A = Adjoint::zero();
    put(A, 0, 0, R); put(A, 0, 3, hat(t) * R); put(A, 3, 3, R);
    return A;
  }
  LIE_HD Mat<T, 4, 4> matrix4() const {
    Mat<T, 4, 4> M = Mat<T, 4, 4>::identity();
    put(M, 0, 0, r.q.matrix()); put_col(M, 0, 3, t);
    return M;
  }
  LIE_HD Mat<T, N, N> projector() const {
    Mat<T, N, N> J = Mat<T, N, N>::zero();
    put(J, 0, 0, M3<T>::identity()); put(J, 0, 3, hat(-t));
    put_quat_projector(J, 3, r.q);
    return J;
  }
  LIE_HD static Adjoint adj(const Tangent& a) {
    V3<T> tau, phi;
    tau[0] = a[0]; tau[1] = a[1]; tau[2] = a[2]; phi[0] = a[3]; phi[1] = a[4]; phi[2] = a[5];
    Adjoint A = Adjoint::zero();
    put(A, 0, 0, hat(phi)); put(A, 0, 3, hat(tau)); put(A, 3, 3, hat(phi));
    return A;
  }
  LIE_HD Tangent Log() const {
    const V3<T> phi = r.Log();
    const V3<T> tau = so3_left_jacobian_inverse(phi) * t;
    Tangent o; o[0] = tau[0]; o[1] = tau[1]; o[2] = tau[2]; o[3] = phi[0]; o[4] = phi[1]; o[5] = phi[2];
    return o;
  }
  LIE_HD static SE3 Exp(const Tangent& a) {
    V3<T> tau, phi;
    tau[0] = a[0]; tau[1] = a[1]; tau[2] = a[2]; phi[0] = a[3]; phi[1] = a[4]; phi[2] = a[5];
    return SE3(SO3<T>::Exp(phi), so3_left_jacobian(phi) * tau);
  }
  // the off-diagonal block of the SE3 left Jacobian (se3.h:146-176)
  LIE_HD static M3<T> calcQ(const Tangent& a) {
    V3<T> tau, phi;
    tau[0] = a[0]; tau[1] = a[1]; tau[2] = a[2]; phi[0] = a[3]; phi[1] = a[4]; phi[2] = a[5];
    const M3<T> Tau = hat(tau), Phi = hat(phi);
    const double th = sqrt((double)phi[0] * phi[0] + (double)phi[1] * phi[1] + (double)phi[2] * phi[2]);
    const double t2 = th * th, t4 = t2 * t2;
    const double c1 = th < kEps ? 1.0 / 6.0 - (1.0 / 120.0) * t2 : (th - sin(th)) / (t2 * th);
    const double c2 = th < kEps ? 1.0 / 24.0 - (1.0 / 720.0) * t2 : (t2 + 2.0 * cos(th) - 2.0) / (2.0 * t4);
    const double c3 = th < kEps ? 1.0 / 120.0 - (1.0 / 2520.0) * t2 : (2.0 * th - 3.0 * sin(th) + th * cos(th)) / (2.0 * t4 * th);
    const M3<T> PT = Phi * Tau, TP = Tau * Phi, PTP = PT * Phi, PP = Phi * Phi;
    return T(0.5) * Tau + T(c1) * (PT + TP + PTP) + T(c2) * (PP * Tau + Tau * PP - T(3) * PTP) + T(c3) * (PTP * Phi + Phi * PTP);
  }
  LIE_HD static Adjoint left_jacobian(const Tangent& a) {
    V3<T> phi; phi[0] = a[3]; phi[1] = a[4]; phi[2] = a[5];
    const M3<T> J = so3_left_jacobian(phi);
    Adjoint A = Adjoint::zero();
    put(A, 0, 0, J); put(A, 0, 3, calcQ(a)); put(A, 3, 3, J);
    return A;
  }
  LIE_HD static Adjoint left_jacobian_inverse(const Tangent& a) {
    V3<T> phi; phi[0] = a[3]; phi[1] = a[4]; phi[2] = a[5];
    const M3<T> Ji = so3_left_jacobian_inverse(phi);
    Adjoint A = Adjoint::zero();
    put(A, 0, 0, Ji); put(A, 0, 3, T(-1) * (Ji * calcQ(a) * Ji)); put(A, 3, 3, Ji);
    return A;
  }
  LIE_HD static Mat<T, 3, K> act_jacobian(const V3<T>& p) {
    Mat<T, 3, K> J;
    put(J, 0, 0, M3<T>::identity()); put(J, 0, 3, hat(-p));
    return J;
  }
  LIE_HD static Mat<T, 4, K> act4_jacobian(const Vec<T, 4>& p) {
    Mat<T, 4, K> J = Mat<T, 4, K>::zero();
    V3<T> p3; p3[0] = p[0]; p3[1] = p[1]; p3[2] = p[2];
    put(J, 0, 0, p[3] * M3<T>::identity()); put(J, 0, 3, hat(-p3));
    return J;
  }
};

// ----------------------------------------------------------------------------------------------------------------- Sim3 ----
template <typename T>
struct Sim3 {
  static constexpr int K = 7, N = 8;
  using Tangent = Vec<T, K>;
  using Adjoint = Mat<T, K, K>;
  V3<T> t;
  RxSO3<T> r;
  LIE_HD Sim3() : t(V3<T>::zero()) {}
  LIE_HD Sim3(const RxSO3<T>& r_, const V3<T>& t_) : t(t_), r(r_) {}
  LIE_HD explicit Sim3(const T* d) : r(d + 3) { t[0] = d[0]; t[1] = d[1]; t[2] = d[2]; }
  LIE_HD void store(T* d) const { d[0] = t[0]; d[1] = t[1]; d[2] = t[2]; r.store(d + 3); }
  LIE_HD Sim3 inv() const { const RxSO3<T> ri = r.inv(); return Sim3(ri, -ri.act(t)); }
  LIE_HD Sim3 operator*(const Sim3& o) const { return Sim3(r * o.r, t + r.act(o.t)); }
  LIE_HD V3<T> act(const V3<T>& p) const { return r.act(p) + t; }
  LIE_HD Vec<T, 4> act4(const Vec<T, 4>& p) const {
    V3<T> p3; p3[0] = p[0]; p3[1] = p[1]; p3[2] = p[2];
    const V3<T> q = r.act(p3) + p[3] * t;
    Vec<T, 4> o; o[0] = q[0]; o[1] = q[1]; o[2] = q[2]; o[3] = p[3];
    return o;
  }
  LIE_HD Mat<T, 4, 4> matrix4() const {
    Mat<T, 4, 4> M = Mat<T, 4, 4>::identity();
    put(M, 0, 0, r.matrix3()); put_col(M, 0, 3, t);
    return M;
  }
  LIE_HD Mat<T, N, N> projector() const {
    Mat<T, N, N> J = Mat<T, N, N>::zero();
    put(J, 0, 0, M3<T>::identity()); put(J, 0, 3, hat(-t)); put_col(J, 0, 6, t);
    put_quat_projector(J, 3, r.q);
    J(7, 6) = r.s;
    return J;
  }
  LIE_HD Adjoint Adj() const {
    Adjoint A = Adjoint::identity();
    const M3<T> R = r.rotation();
    put(A, 0, 0, r.matrix3()); put(A, 0, 3, hat(t) * R); put_col(A, 0, 6, -t); put(A, 3, 3, R);
    return A;
  }
  LIE_HD static Adjoint adj(const Tangent& a) {
    V3<T> tau, phi;
    tau[0] = a[0]; tau[1] = a[1]; tau[2] = a[2]; phi[0] = a[3]; phi[1] = a[4]; phi[2] = a[5];
    Adjoint A = Adjoint::zero();
    put(A, 0, 0, hat(phi) + a[6] * M3<T>::identity()); put(A, 0, 3, hat(tau)); put_col(A, 0, 6, -tau); put(A, 3, 3, hat(phi));
    return A;
  }
  LIE_HD Tangent Log() const {
    const Vec<T, 4> ps = r.Log();
    V3<T> tau = inverse3(RxSO3<T>::calcW(ps)) * t;
    Tangent o; o[0] = tau[0]; o[1] = tau[1]; o[2] = tau[2]; o[3] = ps[0]; o[4] = ps[1]; o[5] = ps[2]; o[6] = ps[3];
    return o;
  }
  LIE_HD static Sim3 Exp(const Tangent& a) {
    V3<T> tau; tau[0] = a[0]; tau[1] = a[1]; tau[2] = a[2];
    Vec<T, 4> ps; ps[0] = a[3]; ps[1] = a[4]; ps[2] = a[5]; ps[3] = a[6];
    return Sim3(RxSO3<T>::Exp(ps), RxSO3<T>::calcW(ps) * tau);
  }
  // series forms (sim3.h:186-211); the reference's sum stops after Xi^4 / 120 (its next line is a separate statement)
  LIE_HD static Adjoint left_jacobian(const Tangent& a) {
    const Adjoint Xi = adj(a), Xi2 = Xi * Xi, Xi4 = Xi2 * Xi2;
    return Adjoint::identity() + T(1.0 / 2.0) * Xi + T(1.0 / 6.0) * Xi2 + T(1.0 / 24.0) * (Xi * Xi2) + T(1.0 / 120.0) * Xi4;
  }
  LIE_HD static Adjoint left_jacobian_inverse(const Tangent& a) {
    const Adjoint Xi = adj(a), Xi2 = Xi * Xi, Xi4 = Xi2 * Xi2;
    return Adjoint::identity() - T(1.0 / 2.0) * Xi + T(1.0 / 12.0) * Xi2 - T(1.0 / 720.0) * Xi4;
  }
  LIE_HD static Mat<T, 3, K> act_jacobian(const V3<T>& p) {
    Mat<T, 3, K> J;
    put(J, 0, 0, M3<T>::identity()); put(J, 0, 3, hat(-p)); put_col(J, 0, 6, p);
    return J;
  }
  LIE_HD static Mat<T, 4, K> act4_jacobian(const Vec<T, 4>& p) {
    Mat<T, 4, K> J = Mat<T, 4, K>::zero();
    V3<T> p3; p3[0] = p[0]; p3[1] = p[1]; p3[2] = p[2];
    put(J, 0, 0, p[3] * M3<T>::identity()); put(J, 0, 3, hat(-p3)); put_col(J, 0, 6, p3);
    return J;
  }
};

}  // namespace lie
