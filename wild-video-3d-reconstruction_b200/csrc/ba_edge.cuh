// Per-edge residual / Jacobian terms shared by the windowed and the global bundle-adjustment kernels.
#pragma once
#include "se3.cuh"
#include "wv_common.cuh"

namespace wv {

struct EdgeTerms {  // both residual rows of one edge
  float Ji[2][6], Jj[2][6], Jz[2], w[2], r[2];
};

// projection, residual, validity mask and Jacobians of one edge (ba_cuda.cu:242-291, 325-331)
template <class ArgsT>
__device__ __forceinline__ void edge_terms(const ArgsT& a, int e, int ix, int jx, float px, float py, float pd,
                                           float fx, float fy, float cx, float cy, EdgeTerms& T) {
  const float* Pi = a.poses + 7 * (size_t)ix;
  const float* Pj = a.poses + 7 * (size_t)jx;
  const Vec3 ti{Pi[0], Pi[1], Pi[2]}, tj{Pj[0], Pj[1], Pj[2]};
  const Quat qi{Pi[3], Pi[4], Pi[5], Pi[6]}, qj{Pj[3], Pj[4], Pj[5], Pj[6]};
  Vec3 tij; Quat qij;
  relative_pose(ti, qi, tj, qj, tij, qij);
  const Vec3 Xi{(px - cx) / fx, (py - cy) / fy, 1.f};
  const Vec3 Rx = rotate(qij, Xi);
  const float W = pd;
  const float X = Rx.x + W * tij.x, Y = Rx.y + W * tij.y, Z = Rx.z + W * tij.z;
  const float d = (Z >= 0.2f) ? 1.f / Z : 0.f;
  const float d2 = d * d;
  const float x1 = fx * (X / Z) + cx, y1 = fy * (Y / Z) + cy;
  const float tx = a.target[2 * (size_t)e], ty = a.target[2 * (size_t)e + 1];
  const float rx = tx - x1, ry = ty - y1;
  const bool ok = (sqrtf(rx * rx + ry * ry) < 128.f) && (Z > 0.2f) && (x1 > -64.f) && (y1 > -64.f) &&
                  (x1 < 2.f * cx + 64.f) && (y1 < 2.f * cy + 64.f);
  const float m = ok ? 1.f : 0.f;
  T.r[0] = rx; T.r[1] = ry;
  T.w[0] = m * a.weight[2 * (size_t)e]; T.w[1] = m * a.weight[2 * (size_t)e + 1];
  T.Jz[0] = fx * (tij.x * d - tij.z * (X * d2));
  T.Jz[1] = fy * (tij.y * d - tij.z * (Y * d2));
  T.Jj[0][0] = fx * W * d; T.Jj[0][1] = 0.f; T.Jj[0][2] = fx * -X * W * d2;
  T.Jj[0][3] = fx * -X * Y * d2; T.Jj[0][4] = fx * (1.f + X * X * d2); T.Jj[0][5] = fx * -Y * d;
  T.Jj[1][0] = 0.f; T.Jj[1][1] = fy * W * d; T.Jj[1][2] = fy * -Y * W * d2;
  T.Jj[1][3] = fy * (-1.f - Y * Y * d2); T.Jj[1][4] = fy * (X * Y * d2); T.Jj[1][5] = fy * X * d;
  adjoint_T(tij, qij, T.Jj[0], T.Ji[0]);
  adjoint_T(tij, qij, T.Jj[1], T.Ji[1]);
}


}  // namespace wv
