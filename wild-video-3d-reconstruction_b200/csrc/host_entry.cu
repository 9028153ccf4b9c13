// Host-pointer entry points of include/wvb200.h: solve_system (CPU, fp64, like the reference) and the
// host-buffer forms of corr / BA that wrap the device entry points with H2D / D2H copies on one stream.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "wv_common.cuh"

namespace {

// The lower triangle of a symmetric matrix in SKYLINE (envelope) storage: row i holds its entries from its first non-zero column
// first[i] to the diagonal, contiguously.  Cholesky fill-in stays inside the envelope, so nothing else is ever stored: a pose
// graph of n frames with a chain and a few loop edges costs O(n * band + loops * n) doubles -- not the (7n)^2 of a dense matrix
// (1.6 GB at n = 2 048) -- where the reference hands the same system to Eigen's sparse SimplicialLLT (dpvo/fastba/ba.cpp:156-172).
struct Skyline {
  std::vector<int64_t> first, off;
  std::vector<double> v;
  void shape(const std::vector<int64_t>& first_col) {
    first = first_col;
    off.resize(first.size() + 1);
    off[0] = 0;
    for (size_t i = 0; i < first.size(); i++) off[i + 1] = off[i] + ((int64_t)i - first[i] + 1);
    v.assign((size_t)off.back(), 0.0);
  }
  double& at(int64_t i, int64_t j) { return v[(size_t)(off[i] + j - first[i])]; }   // first[i] <= j <= i
  const double* row(int64_t i) const { return v.data() + off[i] - first[i]; }      // row(i)[j], first[i] <= j <= i
  double* row(int64_t i) { return v.data() + off[i] - first[i]; }
};

// LL^T in double inside the envelope, then the two substitutions; false if not SPD
bool skyline_cholesky_solve(Skyline& A, std::vector<double>& b) {
  const int64_t n = (int64_t)A.first.size();
  for (int64_t i = 0; i < n; i++) {
    double* ri = A.row(i);
    for (int64_t j = A.first[i]; j <= i; j++) {
      const double* rj = A.row(j);
      double s = ri[j];
      for (int64_t k = std::max(A.first[i], A.first[j]); k < j; k++) s -= ri[k] * rj[k];
      if (j < i) {
        ri[j] = s / rj[j];
      } else {
        if (!(s > 0.0)) return false;
        ri[i] = sqrt(s);
      }
    }
  }
  for (int64_t i = 0; i < n; i++) {
    const double* ri = A.row(i);
    double s = b[i];
    for (int64_t k = A.first[i]; k < i; k++) s -= ri[k] * b[k];
    b[i] = s / ri[i];
  }
  for (int64_t i = n - 1; i >= 0; i--) {
    const double* ri = A.row(i);
    b[i] /= ri[i];
    const double x = b[i];
    for (int64_t k = A.first[i]; k < i; k++) b[k] -= ri[k] * x;
  }
  return true;
}

}  // namespace

extern "C" {

// dpvo/fastba/ba.cpp:174-234: A = J^T J from 7x7 blocks, A_dd += lm * A_dd + ep, b = -J^T res, solve in double.
int wv_solve_system_host(const float* J_i, const float* J_j, const int64_t* ii, const int64_t* jj, const float* res,
                         int64_t r, float ep, float lm, int freen, float* delta, int64_t* n_out) {
  if (r < 0 || (r > 0 && (!J_i || !J_j || !ii || !jj || !res))) return WV_EINVAL;
  int64_t n = 0;
  for (int64_t x = 0; x < r; x++) {
    if (ii[x] < 0 || jj[x] < 0 || ii[x] == jj[x]) return WV_EINVAL;  // the reference exit(1)s on i == j (:205-206)
    n = std::max(n, std::max(ii[x], jj[x]) + 1);
  }
  if (n_out) *n_out = n;
  if (!delta) return WV_OK;
  const int64_t n7 = n * 7;
  int64_t m = freen < 0 ? n7 : std::min<int64_t>(n7, (int64_t)freen * 7);  // solve(A, b, freen*7), ba.cpp:156-172
  for (int64_t a = 0; a < n7; a++) delta[a] = 0.f;
  if (m == 0) return WV_OK;
  // envelope: a row block starts at the lowest block an edge connects it to
  std::vector<int64_t> first((size_t)m);
  for (int64_t a = 0; a < m; a++) first[a] = a / 7 * 7;
  for (int64_t x = 0; x < r; x++) {
    const int64_t lo = std::min(ii[x], jj[x]) * 7, hi = std::max(ii[x], jj[x]) * 7;
    if (hi >= m) continue;
    for (int l = 0; l < 7; l++) first[hi + l] = std::min(first[hi + l], lo);
  }
  Skyline A;
  A.shape(first);
  std::vector<double> b((size_t)m, 0.0);
  for (int64_t x = 0; x < r; x++) {
    const int64_t cb[2] = {ii[x] * 7, jj[x] * 7};
    const float* Jb[2] = {J_i + x * 49, J_j + x * 49};
    for (int p = 0; p < 2; p++) {
      if (cb[p] >= m) continue;
      for (int q = 0; q < 2; q++) {
        if (cb[q] >= m || cb[q] > cb[p]) continue;   // lower triangle only
        for (int l = 0; l < 7; l++)
          for (int c = 0; c < 7; c++) {
            if (cb[q] + c > cb[p] + l) continue;
            double s = 0;
            for (int k = 0; k < 7; k++) s += (double)Jb[p][k * 7 + l] * (double)Jb[q][k * 7 + c];
            A.at(cb[p] + l, cb[q] + c) += s;
          }
      }
      for (int l = 0; l < 7; l++) {
        double s = 0;
        for (int k = 0; k < 7; k++) s += (double)Jb[p][k * 7 + l] * (double)res[x * 7 + k];
        b[(size_t)cb[p] + l] -= s;
      }
    }
  }
  for (int64_t d = 0; d < m; d++) { A.at(d, d) += A.at(d, d) * (double)lm; A.at(d, d) += (double)ep; }
  if (!skyline_cholesky_solve(A, b)) return 1;
  for (int64_t a = 0; a < m; a++) delta[a] = (float)b[a];
  return WV_OK;
}

// ---- host-buffer forms ---------------------------------------------------------------------------
#define HB_TRY(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { rc = ::wv::cuda_fail(_e); goto done; } } while (0)

int wv_corr_forward_host(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii, const int64_t* jj,
                         void* out, int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2,
                         int radius, wv_stream_t stream) {
  if (dtype != WV_F16 && dtype != WV_F32) return WV_EINVAL;
  if ((int64_t)B * E <= 0) return B < 0 || E < 0 ? WV_EINVAL : WV_OK;
  cudaStream_t st = ::wv::as_stream(stream);
  const size_t es = dtype == WV_F16 ? 2 : 4, T = 2 * radius + 1;
  const size_t b1 = (size_t)B * n1 * C * PH * PW * es, b2 = (size_t)B * n2 * C * H2 * W2 * es;
  const size_t bc = (size_t)B * E * 2 * PH * PW * 4, bi = (size_t)E * 8, bo = (size_t)B * E * T * T * PH * PW * es;
  const size_t bw = wv_corr_workspace_bytes(dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius);
  char* d = nullptr;
  int rc = WV_OK, status = 0;
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t o1 = 0, o2 = o1 + al(b1), oc = o2 + al(b2), oi = oc + al(bc), oj = oi + al(bi), oo = oj + al(bi),
               os = oo + al(bo), ow = os + 256, total = ow + al(bw);
  HB_TRY(cudaMallocAsync((void**)&d, total, st));
  HB_TRY(cudaMemcpyAsync(d + o1, fmap1, b1, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + o2, fmap2, b2, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + oc, coords, bc, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + oi, ii, bi, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + oj, jj, bi, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemsetAsync(d + os, 0, 4, st));
  rc = wv_corr_forward(d + o1, d + o2, (const float*)(d + oc), (const int64_t*)(d + oi), (const int64_t*)(d + oj), d + oo,
                       dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius, bw ? d + ow : nullptr, bw, (int*)(d + os), stream);
  if (rc != WV_OK) goto done;
  HB_TRY(cudaMemcpyAsync(out, d + oo, bo, cudaMemcpyDeviceToHost, st));
  HB_TRY(cudaMemcpyAsync(&status, d + os, 4, cudaMemcpyDeviceToHost, st));
  HB_TRY(cudaStreamSynchronize(st));
  if (status & WV_STATUS_BAD_INDEX) rc = WV_EINVAL;
done:
  if (d) cudaFreeAsync(d, st);
  return rc;
}

int wv_ba_forward_host(float* poses, float* patches, const float* intrinsics, const float* target, const float* weight,
                       const float* lmbda, const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                       int64_t n_poses, int64_t n_patches, int P, int t0, int t1, int iterations, const float* patches_est,
                       float mu, int* status_out, wv_stream_t stream) {
  if (E < 0 || n_poses <= 0 || n_patches <= 0 || P <= 0 || t1 < t0) return WV_EINVAL;
  if (status_out) *status_out = 0;
  if (E == 0 || iterations <= 0) return WV_OK;
  cudaStream_t st = ::wv::as_stream(stream);
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t bp = (size_t)n_poses * 7 * 4, bq = (size_t)n_patches * 3 * P * P * 4, bn = (size_t)n_poses * 4 * 4;
  const size_t bt = (size_t)E * 2 * 4, be = (size_t)E * 8, bw = wv_ba_workspace_bytes(E, n_patches, t1 - t0);
  const size_t op = 0, oq = op + al(bp), oe = oq + al(bq), on = oe + (patches_est ? al(bq) : 0), ot = on + al(bn),
               ow = ot + al(bt), ol = ow + al(bt), oi = ol + 256, oj = oi + al(be), ok = oj + al(be), os = ok + al(be),
               owk = os + 256, total = owk + al(bw);
  char* d = nullptr;
  int rc = WV_OK, status = 0;
  HB_TRY(cudaMallocAsync((void**)&d, total, st));
  HB_TRY(cudaMemcpyAsync(d + op, poses, bp, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + oq, patches, bq, cudaMemcpyHostToDevice, st));
  if (patches_est) HB_TRY(cudaMemcpyAsync(d + oe, patches_est, bq, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + on, intrinsics, bn, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + ot, target, bt, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + ow, weight, bt, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + ol, lmbda, 4, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + oi, ii, be, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + oj, jj, be, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemcpyAsync(d + ok, kk, be, cudaMemcpyHostToDevice, st));
  HB_TRY(cudaMemsetAsync(d + os, 0, 4, st));
  rc = wv_ba_forward((float*)(d + op), (float*)(d + oq), (const float*)(d + on), (const float*)(d + ot),
                     (const float*)(d + ow), (const float*)(d + ol), (const int64_t*)(d + oi), (const int64_t*)(d + oj),
                     (const int64_t*)(d + ok), E, n_poses, n_patches, P, t0, t1, iterations,
                     patches_est ? (const float*)(d + oe) : nullptr, mu, d + owk, bw, (int*)(d + os), stream);
  if (rc != WV_OK) goto done;
  HB_TRY(cudaMemcpyAsync(poses, d + op, bp, cudaMemcpyDeviceToHost, st));
  HB_TRY(cudaMemcpyAsync(patches, d + oq, bq, cudaMemcpyDeviceToHost, st));
  HB_TRY(cudaMemcpyAsync(&status, d + os, 4, cudaMemcpyDeviceToHost, st));
  HB_TRY(cudaStreamSynchronize(st));
  if (status_out) *status_out = status;
done:
  if (d) cudaFreeAsync(d, st);
  return rc;
}

}  // extern "C"

// ---- host-buffer session ---------------------------------------------------------------------------
struct wv_session {
  int dev = 0;
  char* arena = nullptr;
  size_t cap = 0;
  cudaStream_t s_in = nullptr, s_cmp = nullptr, s_out = nullptr;
  cudaEvent_t ev_small = nullptr, ev_ba = nullptr, ev_f[4] = {}, ev_c[4] = {};
};

namespace {
__global__ void scale_coords_kernel(const float* __restrict__ in, float* __restrict__ out, float s, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = in[i] * s;
}
}  // namespace

extern "C" {

wv_session* wv_session_create(void) {
  wv_session* s = new wv_session();
  bool ok = cudaGetDevice(&s->dev) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&s->s_in, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&s->s_cmp, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&s->s_out, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&s->ev_small, cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&s->ev_ba, cudaEventDisableTiming) == cudaSuccess;
  for (int l = 0; l < 4 && ok; l++)
    ok = cudaEventCreateWithFlags(&s->ev_f[l], cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&s->ev_c[l], cudaEventDisableTiming) == cudaSuccess;
  if (!ok) { wv_session_destroy(s); return nullptr; }
  return s;
}

void wv_session_destroy(wv_session* s) {
  if (!s) return;
  if (s->arena) cudaFree(s->arena);
  if (s->s_in) cudaStreamDestroy(s->s_in);
  if (s->s_cmp) cudaStreamDestroy(s->s_cmp);
  if (s->s_out) cudaStreamDestroy(s->s_out);
  if (s->ev_small) cudaEventDestroy(s->ev_small);
  if (s->ev_ba) cudaEventDestroy(s->ev_ba);
  for (int l = 0; l < 4; l++) { if (s->ev_f[l]) cudaEventDestroy(s->ev_f[l]); if (s->ev_c[l]) cudaEventDestroy(s->ev_c[l]); }
  delete s;
}

int wv_session_update_host(wv_session* s, wv_update_args* a) {
  if (!s || !a) return WV_EINVAL;
  a->status = 0;
  if ((a->dtype != WV_F16 && a->dtype != WV_F32) || a->nlevels < 0 || a->nlevels > 4 || a->B <= 0 || a->E < 0) return WV_EINVAL;
  if (a->E == 0) return WV_OK;
  const bool do_ba = a->iterations > 0;
  const size_t es = a->dtype == WV_F16 ? 2 : 4, T = 2 * a->radius + 1, PP = (size_t)a->P * a->P;
  auto al = [](size_t x) { return (x + 1023) & ~(size_t)1023; };
  // ---- arena layout
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o += al(bytes); return r; };
  const size_t b1 = (size_t)a->B * a->n1 * a->C * PP * es, bc = (size_t)a->B * a->E * 2 * PP * 4, bi = (size_t)a->E * 8;
  const size_t bo = (size_t)a->B * a->E * T * T * PP * es;
  const size_t o1 = take(b1), oc = take(bc), oii = take(bi), ojj = take(bi), ost = take(256);
  size_t of[4], ocs[4], oo[4], ow[4], bw[4], b2[4];
  for (int l = 0; l < a->nlevels; l++) {
    b2[l] = (size_t)a->B * a->n2 * a->C * a->H2[l] * a->W2[l] * es;
    bw[l] = wv_corr_workspace_bytes(a->dtype, a->B, a->E, a->C, a->P, a->P, a->n1, a->n2, a->H2[l], a->W2[l], a->radius);
    of[l] = take(b2[l]); ocs[l] = take(bc); oo[l] = take(bo); ow[l] = take(bw[l] + 1024);
  }
  // two levels into ONE stacked volume (DPVO.corr's layout) when the caller asks for it and the two-level kernel covers the shape
  const size_t bws = (a->corr_stacked_out && a->nlevels == 2)
                         ? wv_corr_stacked_workspace_bytes(a->dtype, a->B, a->E, a->C, a->P, a->P, a->n1, a->n2, a->H2[0], a->W2[0], a->H2[1], a->W2[1], a->radius)
                         : 0;
  if (a->corr_stacked_out && bws == 0) return WV_EUNSUPPORTED;
  const size_t ows = bws ? take(bws + 1024) : 0, oos = bws ? take(2 * bo) : 0;
  const size_t bp = (size_t)a->n_poses * 7 * 4, bq = (size_t)a->n_patches * 3 * PP * 4, bn = (size_t)a->n_poses * 4 * 4, bt = (size_t)a->E * 2 * 4;
  const size_t bwb = do_ba ? std::max(wv_ba_workspace_bytes(a->E, a->n_patches, a->t1 - a->t0), wv_ba_workspace_bytes(a->E, (int64_t)1 << 22, a->t1 - a->t0)) : 0;
  size_t op = 0, oq = 0, oe = 0, on = 0, otg = 0, owt = 0, ol = 0, oi = 0, oj = 0, ok = 0, owb = 0, onx = 0, onj = 0;
  if (do_ba) {
    op = take(bp); oq = take(bq); oe = a->patches_est ? take(bq) : 0; on = take(bn); otg = take(bt); owt = take(bt); ol = take(256);
    oi = take(bi); oj = take(bi); ok = take(bi); owb = take(bwb + 1024); onx = take(bi); onj = take(bi);
  }
  int rc = WV_OK;
  if (o > s->cap) {
    if (s->arena) { cudaDeviceSynchronize(); cudaFree(s->arena); s->arena = nullptr; s->cap = 0; }
    const size_t want = o + (o >> 3);
    if (cudaMalloc((void**)&s->arena, want) != cudaSuccess) return ::wv::cuda_fail(cudaGetLastError());
    s->cap = want;
  }
  char* d = s->arena;
  int* dstatus = (int*)(d + ost);
#define SS_TRY(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { rc = ::wv::cuda_fail(_e); goto fail; } } while (0)
  // ---- uploads: small pieces first, then the maps from the coarsest level to level 0
  SS_TRY(cudaMemsetAsync(dstatus, 0, 4, s->s_in));
  SS_TRY(cudaMemcpyAsync(d + oc, a->coords, bc, cudaMemcpyHostToDevice, s->s_in));
  SS_TRY(cudaMemcpyAsync(d + oii, a->ii1, bi, cudaMemcpyHostToDevice, s->s_in));
  SS_TRY(cudaMemcpyAsync(d + ojj, a->jj1, bi, cudaMemcpyHostToDevice, s->s_in));
  if (do_ba) {
    SS_TRY(cudaMemcpyAsync(d + op, a->poses, bp, cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaMemcpyAsync(d + oq, a->patches, bq, cudaMemcpyHostToDevice, s->s_in));
    if (a->patches_est) SS_TRY(cudaMemcpyAsync(d + oe, a->patches_est, bq, cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaMemcpyAsync(d + on, a->intrinsics, bn, cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaMemcpyAsync(d + otg, a->target, bt, cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaMemcpyAsync(d + owt, a->weight, bt, cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaMemcpyAsync(d + ol, a->lmbda, 4, cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaMemcpyAsync(d + oi, a->ii, bi, cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaMemcpyAsync(d + oj, a->jj, bi, cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaMemcpyAsync(d + ok, a->kk, bi, cudaMemcpyHostToDevice, s->s_in));
  }
  if (a->nlevels > 0) SS_TRY(cudaMemcpyAsync(d + o1, a->fmap1, b1, cudaMemcpyHostToDevice, s->s_in));
  SS_TRY(cudaEventRecord(s->ev_small, s->s_in));
  for (int l = a->nlevels - 1; l >= 0; l--) {
    SS_TRY(cudaMemcpyAsync(d + of[l], a->fmap2[l], b2[l], cudaMemcpyHostToDevice, s->s_in));
    SS_TRY(cudaEventRecord(s->ev_f[l], s->s_in));
  }
  // ---- compute: BA first (its inputs are there first), then the levels as their maps arrive
  SS_TRY(cudaStreamWaitEvent(s->s_cmp, s->ev_small, 0));
  if (do_ba) {
    if (a->nb_ix && a->nb_jx) {
      rc = wv_neighbors((const int64_t*)(d + ok), (const int64_t*)(d + oj), a->E, (int64_t)1 << 22, (int64_t*)(d + onx), (int64_t*)(d + onj),
                        d + owb, bwb, dstatus, (wv_stream_t)s->s_cmp);
      if (rc != WV_OK) goto fail;
    }
    if (a->nb_ix && a->nb_jx && a->n_patches <= ((int64_t)1 << 22))  // one edge plan serves both calls
      rc = wv_ba_forward_planned((float*)(d + op), (float*)(d + oq), (const float*)(d + on), (const float*)(d + otg),
                                 (const float*)(d + owt), (const float*)(d + ol), (const int64_t*)(d + oi), (const int64_t*)(d + oj),
                                 (const int64_t*)(d + ok), a->E, a->n_poses, a->n_patches, a->P, a->t0, a->t1, a->iterations,
                                 a->patches_est ? (const float*)(d + oe) : nullptr, a->mu, d + owb, bwb, (int64_t)1 << 22, dstatus,
                                 (wv_stream_t)s->s_cmp);
    else
      rc = wv_ba_forward((float*)(d + op), (float*)(d + oq), (const float*)(d + on), (const float*)(d + otg), (const float*)(d + owt),
                         (const float*)(d + ol), (const int64_t*)(d + oi), (const int64_t*)(d + oj), (const int64_t*)(d + ok), a->E,
                         a->n_poses, a->n_patches, a->P, a->t0, a->t1, a->iterations, a->patches_est ? (const float*)(d + oe) : nullptr,
                         a->mu, d + owb, bwb, dstatus, (wv_stream_t)s->s_cmp);
    if (rc != WV_OK) goto fail;
    SS_TRY(cudaEventRecord(s->ev_ba, s->s_cmp));
    SS_TRY(cudaStreamWaitEvent(s->s_out, s->ev_ba, 0));
    SS_TRY(cudaMemcpyAsync(a->poses, d + op, bp, cudaMemcpyDeviceToHost, s->s_out));
    SS_TRY(cudaMemcpyAsync(a->patches, d + oq, bq, cudaMemcpyDeviceToHost, s->s_out));
    if (a->nb_ix && a->nb_jx) {
      SS_TRY(cudaMemcpyAsync(a->nb_ix, d + onx, bi, cudaMemcpyDeviceToHost, s->s_out));
      SS_TRY(cudaMemcpyAsync(a->nb_jx, d + onj, bi, cudaMemcpyDeviceToHost, s->s_out));
    }
  }
  if (bws) {
    SS_TRY(cudaStreamWaitEvent(s->s_cmp, s->ev_f[0], 0));  // level 0 is uploaded last
    char* wsp = (char*)(((uintptr_t)(d + ows) + 1023) & ~(uintptr_t)1023);
    rc = wv_corr_forward_stacked(d + o1, d + of[0], d + of[1], (const float*)(d + oc), a->coord_scale[0], a->coord_scale[1],
                                 (const int64_t*)(d + oii), (const int64_t*)(d + ojj), d + oos, a->dtype, a->B, a->E, a->C, a->P, a->P,
                                 a->n1, a->n2, a->H2[0], a->W2[0], a->H2[1], a->W2[1], a->radius, wsp, bws, dstatus, (wv_stream_t)s->s_cmp);
    if (rc != WV_OK) goto fail;
    SS_TRY(cudaEventRecord(s->ev_c[0], s->s_cmp));
    SS_TRY(cudaStreamWaitEvent(s->s_out, s->ev_c[0], 0));
    SS_TRY(cudaMemcpyAsync(a->corr_stacked_out, d + oos, 2 * bo, cudaMemcpyDeviceToHost, s->s_out));
  }
  for (int l = a->nlevels - 1; l >= 0 && !bws; l--) {
    SS_TRY(cudaStreamWaitEvent(s->s_cmp, s->ev_f[l], 0));
    const float* cl = (const float*)(d + oc);
    if (a->coord_scale[l] != 1.0f) {
      const size_t n = bc / 4;
      scale_coords_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 1184), 256, 0, s->s_cmp>>>((const float*)(d + oc), (float*)(d + ocs[l]), a->coord_scale[l], n);
      cl = (const float*)(d + ocs[l]);
    }
    char* wsp = (char*)(((uintptr_t)(d + ow[l]) + 1023) & ~(uintptr_t)1023);
    rc = wv_corr_forward(d + o1, d + of[l], cl, (const int64_t*)(d + oii), (const int64_t*)(d + ojj), d + oo[l], a->dtype, a->B, a->E, a->C,
                         a->P, a->P, a->n1, a->n2, a->H2[l], a->W2[l], a->radius, bw[l] ? wsp : nullptr, bw[l], dstatus, (wv_stream_t)s->s_cmp);
    if (rc != WV_OK) goto fail;
    SS_TRY(cudaEventRecord(s->ev_c[l], s->s_cmp));
    SS_TRY(cudaStreamWaitEvent(s->s_out, s->ev_c[l], 0));
    SS_TRY(cudaMemcpyAsync(a->corr_out[l], d + oo[l], bo, cudaMemcpyDeviceToHost, s->s_out));
  }
  SS_TRY(cudaStreamSynchronize(s->s_cmp));
  SS_TRY(cudaMemcpyAsync(&a->status, dstatus, 4, cudaMemcpyDeviceToHost, s->s_out));
  SS_TRY(cudaStreamSynchronize(s->s_out));
  if (a->status & WV_STATUS_BAD_INDEX) return WV_EINVAL;
  return WV_OK;
fail:
  cudaStreamSynchronize(s->s_in); cudaStreamSynchronize(s->s_cmp); cudaStreamSynchronize(s->s_out);
  return rc;
#undef SS_TRY
}

}  // extern "C"
