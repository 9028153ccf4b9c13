// Fused projective ops: what dpvo/projective_ops.py builds out of ~12 lietorch / elementwise launches per call
// (`Gij = poses[:, jj] * poses[:, ii].inv()`, a repeat()-materialised broadcast of Gij to E*P*P rows
// (dpvo/lietorch/broadcasting.py:26-29), act4, proj) as ONE kernel per entry point:
//
//   pops_transform   projective_ops.py:53-104 (jacobian=False): coords [E][P][P][2|3] (+ valid [E][P][P])
//   pops_flow_mag    projective_ops.py:111-121: beta * |x_ij - x_ii| + (1 - beta) * |x_ij(t only) - x_ii|      [E][P][P]
//   pops_point_cloud projective_ops.py:106-108: poses[ix]^-1 * iproj(patches, intrinsics[ix])                   [n][P][P][4]
//   frame_distance   dpvo/dpvo.py:383-408 for ALL frame pairs at once: D[i][j] = 0.5 * (mean flow_mag(i -> j over the
//                    M patches of i) + mean flow_mag(j -> i over the M patches of j)), replacing the O(n^2) Python loop with
//                    one `.item()` host sync per pair in get_distance_based_edges (dpvo/dpvo.py:410-429)
//
// Unlike the fastba kernels (row 0 of `intrinsics` for every frame, no depth clamp: ba_cuda.cu:232-238, 411-412) these follow
// projective_ops.py: per-frame intrinsics (`intrinsics[:, ii]` to back-project, `intrinsics[:, jj]` to project) and
// d = 1 / max(Z, 0.1).  The group algebra follows lietorch's construction order (dpvo/lietorch/include/se3.h:36-56, so3.h:31-60):
// a quaternion is re-normalised every time an SO3 is constructed from data.
#include "se3.cuh"
#include "wv_common.cuh"

namespace wv {

__device__ __forceinline__ Quat qnormalize(const Quat& q) {
  const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  return {q.x / n, q.y / n, q.z / n, q.w / n};
}

struct Rigid { Vec3 t; Quat q; };

// poses[j] * poses[i].inv(), as lietorch evaluates it (se3.h:36-38 inv, :45-47 mul; every SO3 constructor normalises)
__device__ __forceinline__ Rigid lie_relative(const float* Pi, const float* Pj) {
  const Quat qi = qnormalize({Pi[3], Pi[4], Pi[5], Pi[6]});
  const Quat qc = qnormalize(conj(qi));                      // so3.inv()
  const Vec3 r = rotate(qc, {Pi[0], Pi[1], Pi[2]});
  const Vec3 tinv{-r.x, -r.y, -r.z};
  const Quat qc2 = qnormalize(qc);                           // the inverse is stored as data and read back by operator*
  const Quat qj = qnormalize({Pj[3], Pj[4], Pj[5], Pj[6]});
  Rigid G;
  G.q = qnormalize(qmul(qj, qc2));
  const Vec3 rj = rotate(qj, tinv);
  G.t = {Pj[0] + rj.x, Pj[1] + rj.y, Pj[2] + rj.z};
  return G;
}

// act4 (se3.h:53-56) on X0 = (xn, yn, 1, d), then proj (projective_ops.py:33-49)
__device__ __forceinline__ void act_project(const Rigid& G, float xn, float yn, float d, float fx, float fy, float cx, float cy,
                                            float& x, float& y, float& dinv, float& Z) {
  const Quat q = qnormalize(G.q);  // SE3(data) on the stored product
  const Vec3 r = rotate(q, {xn, yn, 1.f});
  const float X = r.x + G.t.x * d, Y = r.y + G.t.y * d;
  Z = r.z + G.t.z * d;
  dinv = 1.0f / fmaxf(Z, 0.1f);
  x = fx * (dinv * X) + cx;
  y = fy * (dinv * Y) + cy;
}

struct PopsArgs {
  const float* poses; const float* patches; const float* intr;
  const int64_t* ii; const int64_t* jj; const int64_t* kk;
  int64_t E, n_poses, n_patches;
  int P, tonly, depth;
  float beta;
  float* out; float* valid;
  int* status;
};

// mode 0: transform, mode 1: flow_mag
template <int MODE>
__global__ void pops_edge_kernel(PopsArgs a) {
  const int PP = a.P * a.P;
  const int64_t total = a.E * PP;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = t / PP;
    const int px = (int)(t - e * PP);
    const int64_t i = a.ii[e], j = a.jj[e], k = a.kk[e];
    const int oc = MODE == 0 ? (a.depth ? 3 : 2) : 1;
    float* o = a.out + t * oc;
    if (i < 0 || i >= a.n_poses || j < 0 || j >= a.n_poses || k < 0 || k >= a.n_patches) {
      set_status(a.status, WV_STATUS_BAD_INDEX);
      for (int c = 0; c < oc; c++) o[c] = 0.f;
      if (MODE == 0 && a.valid) a.valid[t] = 0.f;
      continue;
    }
    const float* pc = a.patches + k * 3 * PP + px;
    const float* Ki = a.intr + 4 * i;
    const float* Kj = a.intr + 4 * j;
    const float xn = (pc[0] - Ki[2]) / Ki[0], yn = (pc[PP] - Ki[3]) / Ki[1], d = pc[2 * PP];
    const float* Pi = a.poses + 7 * i;
    const float* Pj = a.poses + 7 * j;
    Rigid G = lie_relative(Pi, Pj);
    float x, y, dinv, Z;
    if (MODE == 0) {
      if (a.tonly) G.q = {0.f, 0.f, 0.f, 1.f};
      act_project(G, xn, yn, d, Kj[0], Kj[1], Kj[2], Kj[3], x, y, dinv, Z);
      o[0] = x; o[1] = y;
      if (a.depth) o[2] = dinv;
      if (a.valid) a.valid[t] = Z > 0.2f ? 1.f : 0.f;
    } else {
      float x0, y0, x2, y2;
      const Rigid G0 = lie_relative(Pi, Pi);  // transform(ii, ii): the reference evaluates the group product, not the identity
      act_project(G0, xn, yn, d, Ki[0], Ki[1], Ki[2], Ki[3], x0, y0, dinv, Z);
      act_project(G, xn, yn, d, Kj[0], Kj[1], Kj[2], Kj[3], x, y, dinv, Z);
      Rigid Gt = G;
      Gt.q = {0.f, 0.f, 0.f, 1.f};
      act_project(Gt, xn, yn, d, Kj[0], Kj[1], Kj[2], Kj[3], x2, y2, dinv, Z);
      const float f1 = sqrtf((x - x0) * (x - x0) + (y - y0) * (y - y0));
      const float f2 = sqrtf((x2 - x0) * (x2 - x0) + (y2 - y0) * (y2 - y0));
      o[0] = a.beta * f1 + (1.f - a.beta) * f2;
    }
  }
}

// transform(..., jacobian=True) (projective_ops.py:71-99): derivatives of the CENTRE pixel's projection with respect to the pose of
// frame j (Jj), of frame i (Ji = -Adj(Gij)^T Jj rows, lietorch adjT) and to the patch's inverse depth (Jz = Jp * Gij.matrix()[:, 3]),
// plus validc = (Z_centre > 0.2).  One thread per edge.
__global__ void pops_jacobian_kernel(PopsArgs a, float* __restrict__ Ji, float* __restrict__ Jj, float* __restrict__ Jz,
                                     float* __restrict__ validc) {
  const int PP = a.P * a.P, c = (a.P / 2) * a.P + a.P / 2;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < a.E; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = a.ii[e], j = a.jj[e], k = a.kk[e];
    float* oi = Ji + e * 12; float* oj = Jj + e * 12; float* oz = Jz + e * 2;
    if (i < 0 || i >= a.n_poses || j < 0 || j >= a.n_poses || k < 0 || k >= a.n_patches) {
      set_status(a.status, WV_STATUS_BAD_INDEX);
      for (int t = 0; t < 12; t++) { oi[t] = 0.f; oj[t] = 0.f; }
      oz[0] = oz[1] = 0.f; validc[e] = 0.f;
      continue;
    }
    const float* pc = a.patches + k * 3 * PP + c;
    const float* Ki = a.intr + 4 * i;
    const float* Kj = a.intr + 4 * j;
    const float xn = (pc[0] - Ki[2]) / Ki[0], yn = (pc[PP] - Ki[3]) / Ki[1], H = pc[2 * PP];
    Rigid G = lie_relative(a.poses + 7 * i, a.poses + 7 * j);
    if (a.tonly) G.q = {0.f, 0.f, 0.f, 1.f};
    const Quat q = qnormalize(G.q);
    const Vec3 r = rotate(q, {xn, yn, 1.f});
    const float X = r.x + G.t.x * H, Y = r.y + G.t.y * H, Z = r.z + G.t.z * H;
    const float d = fabsf(Z) > 0.2f ? 1.0f / Z : 0.f;
    const float fx = Kj[0], fy = Kj[1];
    // Jp (2x4) * Ja (4x6): rows of the projection derivative with respect to a left perturbation of Gj
    const float px0 = fx * d, px2 = -fx * X * d * d, py1 = fy * d, py2 = -fy * Y * d * d;
    float row[2][6] = {{px0 * H, 0.f, px2 * H, px2 * Y, px0 * Z - px2 * X, -px0 * Y},
                       {0.f, py1 * H, py2 * H, -py1 * Z + py2 * Y, -py2 * X, py1 * X}};
#pragma unroll
    for (int rr = 0; rr < 2; rr++) {
      float back[6];
      adjoint_T(G.t, q, row[rr], back);           // Adj(Gij)^T row
#pragma unroll
      for (int t = 0; t < 6; t++) { oj[rr * 6 + t] = row[rr][t]; oi[rr * 6 + t] = -back[t]; }
    }
    oz[0] = px0 * G.t.x + px2 * G.t.z;
    oz[1] = py1 * G.t.y + py2 * G.t.z;
    validc[e] = Z > 0.2f ? 1.f : 0.f;
  }
}

__global__ void pops_point_cloud_kernel(const float* __restrict__ poses, const float* __restrict__ patches,
                                        const float* __restrict__ intr, const int64_t* __restrict__ ix, int64_t n, int64_t n_poses,
                                        int P, float* __restrict__ out, int* status) {
  const int PP = P * P;
  const int64_t total = n * PP;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = t / PP;
    const int px = (int)(t - k * PP);
    const int64_t f = ix[k];
    float* o = out + t * 4;
    if (f < 0 || f >= n_poses) { set_status(status, WV_STATUS_BAD_INDEX); o[0] = o[1] = o[2] = o[3] = 0.f; continue; }
    const float* pc = patches + k * 3 * PP + px;
    const float* K = intr + 4 * f;
    const float xn = (pc[0] - K[2]) / K[0], yn = (pc[PP] - K[3]) / K[1], d = pc[2 * PP];
    const float* Pf = poses + 7 * f;
    const Quat q0 = qnormalize({Pf[3], Pf[4], Pf[5], Pf[6]});
    const Quat qc = qnormalize(conj(q0));
    const Vec3 r = rotate(qc, {Pf[0], Pf[1], Pf[2]});
    const Quat qa = qnormalize(qc);  // SE3(data) of the stored inverse
    const Vec3 R = rotate(qa, {xn, yn, 1.f});
    o[0] = R.x - r.x * d; o[1] = R.y - r.y * d; o[2] = R.z - r.z * d; o[3] = d;
  }
}

// one warp per ordered pair (i, j), i < j: both directions, all M patches x P*P pixels
__global__ void __launch_bounds__(256) frame_distance_kernel(const float* __restrict__ poses, const float* __restrict__ patches,
                                                             const float* __restrict__ intr, int n, int M, int P, float beta,
                                                             float* __restrict__ D) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.y, j = blockIdx.x * 8 + warp;
  if (j >= n) return;
  if (j <= i) {
    if (j == i && lane == 0) D[(size_t)i * n + i] = 0.f;
    return;
  }
  const int PP = P * P, items = M * PP;
  float mean2[2];
#pragma unroll
  for (int dir = 0; dir < 2; dir++) {
    const int s = dir ? j : i, tgt = dir ? i : j;
    const float* Ps = poses + 7 * s;
    const float* Pt = poses + 7 * tgt;
    const float* Ks = intr + 4 * s;
    const float* Kt = intr + 4 * tgt;
    const Rigid G0 = lie_relative(Ps, Ps), G = lie_relative(Ps, Pt);
    Rigid Gt = G;
    Gt.q = {0.f, 0.f, 0.f, 1.f};
    float acc = 0.f;
    for (int t = lane; t < items; t += 32) {
      const int k = t / PP, px = t - k * PP;
      const float* pc = patches + ((size_t)s * M + k) * 3 * PP + px;
      const float xn = (pc[0] - Ks[2]) / Ks[0], yn = (pc[PP] - Ks[3]) / Ks[1], d = pc[2 * PP];
      float x0, y0, x1, y1, x2, y2, dinv, Z;
      act_project(G0, xn, yn, d, Ks[0], Ks[1], Ks[2], Ks[3], x0, y0, dinv, Z);
      act_project(G, xn, yn, d, Kt[0], Kt[1], Kt[2], Kt[3], x1, y1, dinv, Z);
      act_project(Gt, xn, yn, d, Kt[0], Kt[1], Kt[2], Kt[3], x2, y2, dinv, Z);
      acc += beta * sqrtf((x1 - x0) * (x1 - x0) + (y1 - y0) * (y1 - y0)) + (1.f - beta) * sqrtf((x2 - x0) * (x2 - x0) + (y2 - y0) * (y2 - y0));
    }
    mean2[dir] = warp_sum(acc) / (float)items;
  }
  if (lane == 0) {
    const float d = 0.5f * (mean2[0] + mean2[1]);
    D[(size_t)i * n + j] = d;
    D[(size_t)j * n + i] = d;
  }
}

// ---------------------------------------------------------------------------------------------------
static int edge_grid(int64_t total) {
  const int64_t want = (total + 255) / 256, cap = (int64_t)sm_count() * 16;
  return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

int pops_transform(const float* poses, const float* patches, const float* intr, const int64_t* ii, const int64_t* jj, const int64_t* kk,
                   int64_t E, int64_t n_poses, int64_t n_patches, int P, int tonly, int depth, float* coords, float* valid, int* status,
                   cudaStream_t st) {
  if (E == 0) return WV_OK;
  PopsArgs a{poses, patches, intr, ii, jj, kk, E, n_poses, n_patches, P, tonly, depth, 0.f, coords, valid, status};
  pops_edge_kernel<0><<<edge_grid(E * P * P), 256, 0, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int pops_jacobian(const float* poses, const float* patches, const float* intr, const int64_t* ii, const int64_t* jj, const int64_t* kk,
                  int64_t E, int64_t n_poses, int64_t n_patches, int P, int tonly, float* Ji, float* Jj, float* Jz, float* validc, int* status,
                  cudaStream_t st) {
  if (E == 0) return WV_OK;
  PopsArgs a{poses, patches, intr, ii, jj, kk, E, n_poses, n_patches, P, tonly, 0, 0.f, nullptr, nullptr, status};
  pops_jacobian_kernel<<<edge_grid(E), 256, 0, st>>>(a, Ji, Jj, Jz, validc);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int pops_flow_mag(const float* poses, const float* patches, const float* intr, const int64_t* ii, const int64_t* jj, const int64_t* kk,
                  int64_t E, int64_t n_poses, int64_t n_patches, int P, float beta, float* out, int* status, cudaStream_t st) {
  if (E == 0) return WV_OK;
  PopsArgs a{poses, patches, intr, ii, jj, kk, E, n_poses, n_patches, P, 0, 0, beta, out, nullptr, status};
  pops_edge_kernel<1><<<edge_grid(E * P * P), 256, 0, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int pops_point_cloud(const float* poses, const float* patches, const float* intr, const int64_t* ix, int64_t n, int64_t n_poses, int P,
                     float* out, int* status, cudaStream_t st) {
  if (n == 0) return WV_OK;
  pops_point_cloud_kernel<<<edge_grid(n * P * P), 256, 0, st>>>(poses, patches, intr, ix, n, n_poses, P, out, status);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int frame_distance(const float* poses, const float* patches, const float* intr, int n, int M, int P, float beta, float* D,
                   cudaStream_t st) {
  if (n == 0) return WV_OK;
  frame_distance_kernel<<<dim3((n + 7) / 8, n), 256, 0, st>>>(poses, patches, intr, n, M, P, beta, D);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

}  // namespace wv
