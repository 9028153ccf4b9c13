/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of the DPVO altcorr + fastba hot path.
 *
 * Plain C, no CUDA, no torch.  Each function cites the reference file:line it follows
 * (paths under /root/reference).  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may call into this library; the product
 * (wild-video-3d-reconstruction_b200/) never links, imports or executes it.
 *
 * PARITY PIN: the reference ships no golden vectors for this path.  The oracle
 * is pinned instead against the reference ITSELF: oracle/build_ref.py compiles the reference's
 * own correlation_kernel.cu / ba_cuda.cu / ba.cpp / lietorch_cpu.cpp as host code into
 * oracle/_ref/, tests/test_oracle_vs_ref.py compares the two on seeded inputs, and
 * oracle/make_golden.py freezes reference outputs as fixtures under tests/golden/.
 *
 * precision modes (argument `mode`):
 *   ORC_F64    everything in double: the "truth" used for error measurements
 *   ORC_F32    float arithmetic in the reference's evaluation order (edges / channels in
 *              index order, no FMA contraction)
 *   ORC_F16EMU corr only: emulates the reference's scalar_t=half kernel (products and the
 *              running sum rounded to half after every step, blend in half)
 */
#ifndef WV_ORACLE_H_
#define WV_ORACLE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_F64 = 0, ORC_F32 = 1, ORC_F16EMU = 2 };

/* correlation_kernel.cu:82-136 (corr_forward_kernel) + :193-233 (host bilinear blend + permute).
 * fmap1 [B][n1][C][PH][PW], fmap2 [B][n2][C][H2][W2] as float (fp16 data widened exactly),
 * coords [B][E][2][PH][PW], ii/jj [E].  out [B][E][D-1][D-1][PH][PW] (dim2 = x-tap, dim3 = y-tap),
 * D = 2*radius+2; out is double for ORC_F64, float otherwise.  returns 0, or -1 on a bad index. */
int orc_corr_forward(const float* fmap1, const float* fmap2, const float* coords,
                     const int64_t* ii, const int64_t* jj, void* out,
                     int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2,
                     int radius, int mode, int nthreads);

/* correlation_kernel.cu:139-190 + :236-286.  grad [B][E][D-1][D-1][PH][PW] (float), outputs are
 * double [like fmap1], [like fmap2], accumulated in double (truth only). */
int orc_corr_backward(const float* fmap1, const float* fmap2, const float* coords,
                      const int64_t* ii, const int64_t* jj, const float* grad,
                      double* fmap1_grad, double* fmap2_grad,
                      int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int radius);

/* correlation_kernel.cu:16-47, 288-310.  net [B][C][H][W], coords [B][M][2] -> patches [B][M][C][D][D] (zero filled) */
int orc_patchify_forward(const float* net, const float* coords, float* patches,
                         int B, int M, int C, int H, int W, int radius);
/* correlation_kernel.cu:49-80, 313-333.  gradient [B][M][C][D][D] -> net_grad [B][C][H][W] (double) */
int orc_patchify_backward(const float* gradient, const float* coords, double* net_grad,
                          int B, int M, int C, int H, int W, int radius);

/* ba_cuda.cu:422-540 (cuda_ba driver) with :214-365 (edge kernel), :160-211 (retractions), :18-156 (SE3 helpers).
 * poses [n_poses][7], patches [n_patches][3][P][P], intrinsics [n_poses][4] (row 0 used), target/weight [E][2].
 * In-place on poses/patches (float storage; ORC_F64 widens, iterates in double, narrows at the end).
 * patches_est (nullable) + mu: depth-prior term of dpvo/ba.py:151-159 added to C and u.
 * dX_out (nullable) [iterations][t1-t0][6], dZ_out (nullable) [iterations][Mp]: per-iteration updates (double).
 * returns 0, 1 when the reduced system is not positive definite (reference: C++ exception), -1 bad index. */
int orc_ba_forward(float* poses, float* patches, const float* intrinsics,
                   const float* target, const float* weight, const float* lmbda,
                   const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                   int64_t n_poses, int64_t n_patches, int P, int t0, int t1, int iterations,
                   const float* patches_est, float mu, int mode,
                   double* dX_out, double* dZ_out, int64_t* Mp_out);

/* ba_cuda.cu:368-418, 543-575.  coords [E][2][P][P] */
int orc_reproject(const float* poses, const float* patches, const float* intrinsics,
                  const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                  int64_t n_poses, int64_t n_patches, int P, float* coords, int mode);

/* ba.cpp:106-151.  for each edge, index of the previous / next edge of the same `ii` group in
 * (jj, original index) order, -1 at the ends. */
int orc_neighbors(const int64_t* ii, const int64_t* jj, int64_t E, int64_t* ix, int64_t* jx);

/* ba.cpp:174-234 (solve_system) + :156-172 (solve): PGO normal equations from 7x7 blocks, LM damping,
 * double Cholesky (dense here; the reference uses Eigen::SimplicialCholesky<double>, same solution).
 * J_i, J_j [r][7][7], res [r][7] -> delta [n][7], n = max(ii,jj)+1 (written to *n_out; delta must hold it).
 * freen < 0: full solve; else only the top-left freen*7 block is solved, rest zero.  returns 0, 1 if not SPD. */
int orc_solve_system(const float* J_i, const float* J_j, const int64_t* ii, const int64_t* jj,
                     const float* res, int64_t r, float ep, float lm, int freen,
                     float* delta, int64_t* n_out);

#ifdef __cplusplus
}
#endif
#endif
