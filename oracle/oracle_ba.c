/* TEST INFRASTRUCTURE ONLY -- see wv_oracle.h.  CPU restatement of dpvo/fastba. */
#include "wv_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define REAL float
#define FN(x) x##_f32
#define SQRT sqrtf
#define SIN sinf
#define COS cosf
#include "oracle_ba_impl.inc"
#undef REAL
#undef FN
#undef SQRT
#undef SIN
#undef COS

#define REAL double
#define FN(x) x##_f64
#define SQRT sqrt
#define SIN sin
#define COS cos
#include "oracle_ba_impl.inc"
#undef REAL
#undef FN
#undef SQRT
#undef SIN
#undef COS

/* torch::_unique(kk, sorted=true, return_inverse=true), ba_cuda.cu:435-437 */
static int64_t unique_sorted(const int64_t* kk, int64_t E, int64_t n_patches, int64_t* ku, int64_t** kx_out) {
  int64_t* rank = (int64_t*)calloc((size_t)n_patches + 1, sizeof(int64_t));
  for (int64_t e = 0; e < E; e++) rank[kk[e]] = 1;
  int64_t Mp = 0;
  for (int64_t k = 0; k < n_patches; k++) Mp += rank[k];
  int64_t* kx = (int64_t*)malloc(sizeof(int64_t) * (size_t)(Mp + 1));
  int64_t c = 0;
  for (int64_t k = 0; k < n_patches; k++)
    if (rank[k]) { kx[c] = k; rank[k] = c++; }
  for (int64_t e = 0; e < E; e++) ku[e] = rank[kk[e]];
  free(rank);
  *kx_out = kx;
  return Mp;
}

static int check_edges(const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                       int64_t n_poses, int64_t n_patches) {
  for (int64_t e = 0; e < E; e++)
    if (ii[e] < 0 || ii[e] >= n_poses || jj[e] < 0 || jj[e] >= n_poses || kk[e] < 0 || kk[e] >= n_patches)
      return -1;
  return 0;
}

int orc_ba_forward(float* poses, float* patches, const float* intrinsics,
                   const float* target, const float* weight, const float* lmbda,
                   const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                   int64_t n_poses, int64_t n_patches, int P, int t0, int t1, int iterations,
                   const float* patches_est, float mu, int mode,
                   double* dX_out, double* dZ_out, int64_t* Mp_out) {
  if (check_edges(ii, jj, kk, E, n_poses, n_patches)) return -1;
  if (t1 < t0 || t0 < 0 || t1 > n_poses) return -1;
  int64_t* ku = (int64_t*)malloc(sizeof(int64_t) * (size_t)(E + 1));
  int64_t* kx = NULL;
  const int64_t Mp = unique_sorted(kk, E, n_patches, ku, &kx);
  if (Mp_out) *Mp_out = Mp;
  int rc;
  if (mode == ORC_F32) {
    rc = ba_run_f32(poses, patches, intrinsics, target, weight, lmbda[0], ii, jj, kk, ku, kx, E, Mp, P,
                    t0, t1, iterations, patches_est, mu, dX_out, dZ_out);
  } else {
    const size_t np = (size_t)n_poses * 7, nq = (size_t)n_patches * 3 * P * P;
    double* dp = (double*)malloc(sizeof(double) * np);
    double* dq = (double*)malloc(sizeof(double) * nq);
    double* de = patches_est ? (double*)malloc(sizeof(double) * nq) : NULL;
    double* dt = (double*)malloc(sizeof(double) * (size_t)(2 * E + 1));
    double* dw = (double*)malloc(sizeof(double) * (size_t)(2 * E + 1));
    double di[4] = {intrinsics[0], intrinsics[1], intrinsics[2], intrinsics[3]};
    for (size_t i = 0; i < np; i++) dp[i] = poses[i];
    for (size_t i = 0; i < nq; i++) dq[i] = patches[i];
    if (de) for (size_t i = 0; i < nq; i++) de[i] = patches_est[i];
    for (int64_t i = 0; i < 2 * E; i++) { dt[i] = target[i]; dw[i] = weight[i]; }
    rc = ba_run_f64(dp, dq, di, dt, dw, (double)lmbda[0], ii, jj, kk, ku, kx, E, Mp, P, t0, t1, iterations,
                    de, (double)mu, dX_out, dZ_out);
    for (size_t i = 0; i < np; i++) poses[i] = (float)dp[i];
    for (size_t i = 0; i < nq; i++) patches[i] = (float)dq[i];
    free(dp); free(dq); free(de); free(dt); free(dw);
  }
  free(ku); free(kx);
  return rc;
}

int orc_reproject(const float* poses, const float* patches, const float* intrinsics,
                  const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                  int64_t n_poses, int64_t n_patches, int P, float* coords, int mode) {
  if (check_edges(ii, jj, kk, E, n_poses, n_patches)) return -1;
  if (mode == ORC_F32) {
    reproject_run_f32(poses, patches, intrinsics, ii, jj, kk, E, P, coords);
    return 0;
  }
  const size_t np = (size_t)n_poses * 7, nq = (size_t)n_patches * 3 * P * P, nc = (size_t)E * 2 * P * P;
  double* dp = (double*)malloc(sizeof(double) * np);
  double* dq = (double*)malloc(sizeof(double) * nq);
  double* dc = (double*)malloc(sizeof(double) * (nc + 1));
  double di[4] = {intrinsics[0], intrinsics[1], intrinsics[2], intrinsics[3]};
  for (size_t i = 0; i < np; i++) dp[i] = poses[i];
  for (size_t i = 0; i < nq; i++) dq[i] = patches[i];
  reproject_run_f64(dp, dq, di, ii, jj, kk, E, P, dc);
  for (size_t i = 0; i < nc; i++) coords[i] = (float)dc[i];
  free(dp); free(dq); free(dc);
  return 0;
}

/* ---- neighbors: ba.cpp:106-151 ---------------------------------------------------------- */
typedef struct { int64_t g, j, i; } nb_key;
static int nb_cmp(const void* a, const void* b) {
  const nb_key* x = (const nb_key*)a; const nb_key* y = (const nb_key*)b;
  if (x->g != y->g) return x->g < y->g ? -1 : 1;
  if (x->j != y->j) return x->j < y->j ? -1 : 1; /* std::stable_sort by jj over idx lists built in */
  return x->i < y->i ? -1 : (x->i > y->i);       /* increasing i  ==  ties broken by original index */
}

int orc_neighbors(const int64_t* ii, const int64_t* jj, int64_t E, int64_t* ix, int64_t* jx) {
  nb_key* k = (nb_key*)malloc(sizeof(nb_key) * (size_t)(E + 1));
  for (int64_t e = 0; e < E; e++) { k[e].g = ii[e]; k[e].j = jj[e]; k[e].i = e; }
  qsort(k, (size_t)E, sizeof(nb_key), nb_cmp);
  for (int64_t r = 0; r < E; r++) {
    ix[k[r].i] = (r > 0 && k[r - 1].g == k[r].g) ? k[r - 1].i : -1;
    jx[k[r].i] = (r + 1 < E && k[r + 1].g == k[r].g) ? k[r + 1].i : -1;
  }
  free(k);
  return 0;
}

/* ---- solve_system: ba.cpp:174-234 ----------------------------------------------------------- */
int orc_solve_system(const float* J_i, const float* J_j, const int64_t* ii, const int64_t* jj,
                     const float* res, int64_t r, float ep, float lm, int freen,
                     float* delta, int64_t* n_out) {
  int64_t n = 0;
  for (int64_t x = 0; x < r; x++) {
    if (ii[x] == jj[x] || ii[x] < 0 || jj[x] < 0) return -1; /* reference: exit(1) on i == j, :205-206 */
    if (ii[x] + 1 > n) n = ii[x] + 1;
    if (jj[x] + 1 > n) n = jj[x] + 1;
  }
  if (n_out) *n_out = n;
  const int64_t n7 = n * 7;
  double* A = (double*)calloc((size_t)(n7 * n7 + 1), sizeof(double));
  double* b = (double*)calloc((size_t)n7 + 1, sizeof(double));
  /* J is (r*7) x (n*7) with blocks J_i at column block i, J_j at column block j (:200-214);
   * values are cast to double as Eigen::Triplet<double> does.  A = J^T J, b = -J^T res. */
  for (int64_t x = 0; x < r; x++) {
    const int64_t cb[2] = {ii[x] * 7, jj[x] * 7};
    const float* Jb[2] = {J_i + x * 49, J_j + x * 49};
    for (int p = 0; p < 2; p++)
      for (int q = 0; q < 2; q++)
        for (int l = 0; l < 7; l++)
          for (int m = 0; m < 7; m++) {
            double s = 0;
            for (int k = 0; k < 7; k++) s += (double)Jb[p][k * 7 + l] * (double)Jb[q][k * 7 + m];
            A[(cb[p] + l) * n7 + cb[q] + m] += s;
          }
    for (int p = 0; p < 2; p++)
      for (int l = 0; l < 7; l++) {
        double s = 0;
        for (int k = 0; k < 7; k++) s += (double)Jb[p][k * 7 + l] * (double)res[x * 7 + k];
        b[cb[p] + l] -= s;
      }
  }
  for (int64_t d = 0; d < n7; d++) { A[d * n7 + d] += A[d * n7 + d] * (double)lm; A[d * n7 + d] += (double)ep; } /* :221-222 */
  int64_t m = (freen < 0) ? n7 : (int64_t)freen * 7; /* solve(A, b, freen*7), :156-172 */
  if (m > n7) m = n7;
  double* As = (double*)malloc(sizeof(double) * (size_t)(m * m + 1));
  double* bs = (double*)malloc(sizeof(double) * (size_t)(m + 1));
  for (int64_t a = 0; a < m; a++) { bs[a] = b[a]; for (int64_t c = 0; c < m; c++) As[a * m + c] = A[a * n7 + c]; }
  const int rc = chol_solve_f64(As, bs, (int)m);
  for (int64_t a = 0; a < n7; a++) delta[a] = (a < m && rc == 0) ? (float)bs[a] : 0.f;
  free(A); free(b); free(As); free(bs);
  return rc;
}
