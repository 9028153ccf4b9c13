/* wvb200.h -- C ABI of the B200-native (sm_100a) altcorr + fastba hot path.
 *
 * Drop-in boundary for ljjTYJR/Wild-video-3d-reconstruction's DPVO tracker: these entry points are what
 * the reference's two extension modules bind (citations are paths under the reference tree):
 *
 *   cuda_corr.forward / backward / patchify_forward / patchify_backward   dpvo/altcorr/correlation.cpp:57-63
 *   cuda_ba.forward / neighbors / reproject / solve_system                dpvo/fastba/ba.cpp:236-241
 *   lietorch_backends.{expm, logm, inv, mul, adj, adjT, act, act4}[_backward], as_matrix, projector, Jinv
 *                                                                          dpvo/lietorch/src/lietorch.cpp:286-316
 *
 * Conventions
 *   - plain pointers and sizes only; no torch types.  Unless a function says "host", every pointer is a
 *     DEVICE pointer on the current CUDA device and work is enqueued on `stream` without host
 *     synchronisation (capture-safe).  The caller owns all buffers.
 *   - return value: WV_OK (0) or a negative WV_E* code; wv_strerror() names it.  No exceptions cross
 *     the ABI.  Data-dependent failures that only the device can see (an edge index out of range,
 *     a reduced camera system that is not positive definite) are reported through a caller-supplied
 *     device word `status` (see WV_STATUS_*), which the caller reads back when it wants the
 *     reference's "raise in the same call" behaviour.
 *   - dtype of feature maps: WV_F16 or WV_F32 (the tracker runs fp16, training fp32).  Geometry is fp32,
 *     indices are int64 -- exactly the accessor types of the reference kernels.
 */
#ifndef WVB200_H_
#define WVB200_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* wv_stream_t; /* == cudaStream_t */

enum { WV_F16 = 0, WV_F32 = 1, WV_F64 = 2 /* lietorch only */ };

enum {
  WV_OK = 0,
  WV_EINVAL = -1,      /* bad argument (null pointer, negative size, unsupported dtype/shape) */
  WV_ECUDA = -2,       /* a CUDA runtime/driver call failed; wv_last_cuda_error() has the code */
  WV_EWORKSPACE = -3,  /* workspace too small */
  WV_EUNSUPPORTED = -4, /* shape outside what this build supports */
  WV_ELIMIT = -5        /* a size exceeds the fixed maximum of a no-allocation entry point (wv_neighbors_static) */
};

/* bits of the device status word */
enum {
  WV_STATUS_BAD_INDEX = 1, /* an ii/jj/kk entry was outside its tensor; that edge was skipped */
  WV_STATUS_NOT_SPD = 2,   /* Cholesky met a non-positive pivot (reference: torch::linalg::cholesky throws, ba_cuda.cu:521) */
  WV_STATUS_TOO_DENSE = 4, /* global BA: one patch had more than 8 256 entries (two per pose of a 4 096-frame buffer) in its column of E; the excess was dropped */
  WV_STATUS_STALE_PLAN = 8 /* wv_ba_forward_planned: the plan in the workspace was not built for this kk; NOTHING was applied (poses / patches untouched): call wv_ba_forward */
};

const char* wv_strerror(int code);
int wv_last_cuda_error(void);
/* number of kernels this library has launched in this process so far (every launch site counts itself); bench.py reports the
 * difference over its timed region as "gpu_launches" */
unsigned long long wv_launch_count(void);
/* measurement aid: enqueue an empty cooperative grid of `n_ctas` CTAs (<= 0: the BA kernel's grid) that executes `iters`
 * grid-wide barriers; timing it gives the per-barrier latency the windowed BA's latency model uses (bench.py roofline_ba) */
int wv_gridsync_probe(int n_ctas, int iters, wv_stream_t stream);
/* 1 when a device of compute capability 10.x is current, else 0 (the library has sm_100a code only) */
int wv_device_supported(void);
/* The library's own status word of the CURRENT device (allocated and zeroed on first use, never freed): one word shared by
 * every module bound to this library, so that an error flagged by one (e.g. a bad index in a correlation call) is seen by
 * whoever reads the word next.  Callers may pass their own word instead wherever a `status` argument is taken.  NULL on failure.
 * No reference counterpart (the reference raises C++ exceptions). */
int* wv_status_word(void);
/* Read a device status word back to the host after everything enqueued on `stream` so far, and clear it when non-zero.
 * Blocks (a short spin on an event, then a blocking wait); *value receives the WV_STATUS_* bits.  This is the one host
 * synchronisation a caller needs to turn a data-dependent failure into an exception from the same call, as the reference's
 * torch::linalg::cholesky does (dpvo/fastba/ba_cuda.cu:521). */
int wv_status_read(int* status, int* value, wv_stream_t stream);
/* The same read for 1 <= n <= 16 consecutive device ints that the host needs to size an output (the group count of
 * wv_group_plan); never clears. */
int wv_words_read(const int* words, int n, int* values, wv_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * altcorr
 * ------------------------------------------------------------------------------------------- */

/* scratch bytes wv_corr_forward wants for its channels-last staging of the referenced fmap2 slots
 * (0 when the generic path will be taken).  Reference: none (it gathers NCHW directly). */
size_t wv_corr_workspace_bytes(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int radius);

/* cuda_corr.forward -- replaces corr_cuda_forward + corr_forward_kernel + the ATen bilinear blend
 * (dpvo/altcorr/correlation_kernel.cu:82-136, 193-233).
 *   fmap1 [B][n1][C][PH][PW], fmap2 [B][n2][C][H2][W2] (dtype), coords [B][E][2][PH][PW] fp32,
 *   ii [E] in [0,n1), jj [E] in [0,n2)  ->  out [B][E][2r+1][2r+1][PH][PW] (dtype), contiguous in the
 *   reference's *permuted* order (dim2 = x-tap, dim3 = y-tap; correlation_kernel.cu:232).
 *   Products are accumulated in fp32 for both dtypes (the reference accumulates in scalar_t).
 *   workspace may be NULL (generic path).  status may be NULL. */
int wv_corr_forward(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii,
                    const int64_t* jj, void* out, int dtype, int B, int E, int C, int PH, int PW,
                    int n1, int n2, int H2, int W2, int radius, void* workspace, size_t workspace_bytes,
                    int* status, wv_stream_t stream);

/* DPVO.corr (dpvo/dpvo.py:325-333) as ONE call: both pyramid levels of every edge and the final
 * torch.stack([corr1, corr2], -1).view(1, E, -1), replacing 2 x wv_corr_forward + the stack (a read of both
 * volumes and a write of the stacked one).  Tracker configuration only (WV_F16, C = 128, 3x3 patches, radius 3):
 * wv_corr_stacked_workspace_bytes returns 0 for anything else and the caller composes the two-call form.
 *   fmap2_l0 [B][n2][C][H0][W0], fmap2_l1 [B][n2][C][H1][W1] (the same ring slots jj at both levels),
 *   level l samples at coords * cmul_l (the reference divides by 1 and 4: pass 1.0f and 0.25f -- exact),
 *   out [B][E][2r+1][2r+1][PH][PW][2] (dtype): [...][0] = level 0, [...][1] = level 1.
 * workspace is required.  status may be NULL. */
size_t wv_corr_stacked_workspace_bytes(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2,
                                       int H0, int W0, int H1, int W1, int radius);
int wv_corr_forward_stacked(const void* fmap1, const void* fmap2_l0, const void* fmap2_l1,
                            const float* coords, float cmul0, float cmul1, const int64_t* ii,
                            const int64_t* jj, void* out, int dtype, int B, int E, int C, int PH, int PW,
                            int n1, int n2, int H0, int W0, int H1, int W1, int radius, void* workspace,
                            size_t workspace_bytes, int* status, wv_stream_t stream);

/* cuda_corr.backward -- replaces corr_cuda_backward + corr_backward_kernel (correlation_kernel.cu:139-190, 236-286).
 *   grad [B][E][2r+1][2r+1][PH][PW] fp32 (same permuted order as `out`), fmap1_grad / fmap2_grad fp32
 *   accumulators shaped like fmap1 / fmap2, ZEROED BY THE CALLER; the binding casts them to dtype. */
int wv_corr_backward(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii,
                     const int64_t* jj, const float* grad, float* fmap1_grad, float* fmap2_grad,
                     int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2,
                     int radius, int* status, wv_stream_t stream);

/* cuda_corr.patchify_forward (correlation_kernel.cu:16-47, 288-310): net [B][C][H][W], coords [B][M][2]
 * -> patches [B][M][C][D][D], D = 2r+2, zero where the tap is outside the map (written by this call). */
int wv_patchify_forward(const void* net, const float* coords, void* patches, int dtype, int B, int M,
                        int C, int H, int W, int radius, wv_stream_t stream);

/* cuda_corr.patchify_backward (correlation_kernel.cu:49-80, 313-333): gradient [B][M][C][D][D] (dtype)
 * -> net_grad [B][C][H][W] fp32 accumulator ZEROED BY THE CALLER. */
int wv_patchify_backward(const void* gradient, const float* coords, float* net_grad, int dtype, int B,
                         int M, int C, int H, int W, int radius, wv_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * fastba
 * ------------------------------------------------------------------------------------------- */

/* workspace for wv_ba_forward / wv_neighbors on E edges over a patch buffer of n_patches rows */
size_t wv_ba_workspace_bytes(int64_t E, int64_t n_patches, int n_free_poses);

/* cuda_ba.forward -- replaces cuda_ba() (dpvo/fastba/ba_cuda.cu:422-540): torch::_unique(kk), the edge
 * kernel reprojection_residuals_and_hessian (:214-365), the ATen Schur/Cholesky sequence (:492-526) and
 * both retraction kernels (:160-211), for `iterations` Gauss-Newton steps, IN PLACE on poses / patches.
 *   poses [n_poses][7] (tx,ty,tz,qx,qy,qz,qw; world->camera), patches [n_patches][3][P][P],
 *   intrinsics [n_poses][4] (row 0 is used, ba_cuda.cu:232-238), target / weight [E][2], lmbda [1],
 *   ii / jj in [0,n_poses), kk in [0,n_patches); poses t0..t1-1 are free, t1 == t0 is structure-only.
 *   patches_est (nullable) [n_patches][3][P][P] + mu: the depth-prior term of dpvo/ba.py:151-159
 *   (C_k += mu*L_k, u_k -= mu*L_k*(d_k - dhat_k), L_k = dhat_k > 0); NULL reproduces the 12-argument call.
 *   status: device word, OR-ed with WV_STATUS_*; on WV_STATUS_NOT_SPD the failing iteration and all later
 *   ones apply no update (the reference throws before its retraction kernels). */
int wv_ba_forward(float* poses, float* patches, const float* intrinsics, const float* target,
                  const float* weight, const float* lmbda, const int64_t* ii, const int64_t* jj,
                  const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0, int t1,
                  int iterations, const float* patches_est, float mu, void* workspace,
                  size_t workspace_bytes, int* status, wv_stream_t stream);

/* wv_ba_forward when `workspace` already holds the edge plan that wv_neighbors(kk, jj, E, n_groups = plan_groups, ...)
 * built for the SAME kk / jj (the tracker calls neighbors and BA on one edge list per update(), dpvo/net.py:81 then
 * dpvo/dpvo.py:733): skips the second plan build.  workspace_bytes >= wv_ba_workspace_bytes(E, plan_groups, t1 - t0);
 * kk values must be < plan_groups.  ii / jj are range-checked by the BA kernel itself, every edge's kk is verified against the
 * plan (mismatch -> WV_STATUS_STALE_PLAN: the kernel stops before it applies anything, the caller repeats with wv_ba_forward), and patch ids >= n_patches are flagged
 * (WV_STATUS_BAD_INDEX) and skipped. */
int wv_ba_forward_planned(float* poses, float* patches, const float* intrinsics, const float* target,
                          const float* weight, const float* lmbda, const int64_t* ii, const int64_t* jj,
                          const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0,
                          int t1, int iterations, const float* patches_est, float mu, void* workspace,
                          size_t workspace_bytes, int64_t plan_groups, int* status, wv_stream_t stream);

/* ---- BA with more than 16 free poses (the loop-closure / end-of-run global BA, dpvo/dpvo.py:436-505) ----------
 * The same cuda_ba() (dpvo/fastba/ba_cuda.cu:422-540), split at its two natural seams so that (a) the dense
 * 6N x 6N solve can be done by the caller's solver (the torch binding uses the reference's own
 * torch::linalg::cholesky + cholesky_solve, ba_cuda.cu:521-523) and (b) the patches can be sharded over GPUs:
 * rank `shard` of `nshards` accumulates only its contiguous run of the sorted unique patch ids, the caller sums
 * [S | y] over ranks (one allreduce per iteration), solves, and sums dZ over ranks.  One iteration is
 *     wv_ba_global_build -> (allreduce S, y) -> wv_ba_global_damp -> dX = S^-1 y (caller) ->
 *     wv_ba_global_dz -> (allreduce dZ) -> wv_ba_global_retract.
 * The dense 6N x Mp matrix E of the reference (:443-444) is never formed.
 *   S [6N][6N] row-major, 16-byte aligned: only the LOWER triangle is written (the upper is left zero); y [6N]; N = t1 - t0.
 *   dX [N][6]; dZ [min(E, n_patches)] indexed by the rank of the patch id among the sorted unique kk.
 *   build_plan != 0 rebuilds the edge plan in `workspace` (needed once per edge list, not per iteration).
 *   workspace_bytes >= wv_ba_global_workspace_bytes(E, n_patches); the same workspace goes to all four calls. */
size_t wv_ba_global_workspace_bytes(int64_t E, int64_t n_patches);
int wv_ba_global_build(float* poses, float* patches, const float* intrinsics, const float* target,
                       const float* weight, const float* lmbda, const int64_t* ii, const int64_t* jj,
                       const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0, int t1,
                       const float* patches_est, float mu, int shard, int nshards, int build_plan, float* S,
                       float* y, void* workspace, size_t workspace_bytes, int* status, wv_stream_t stream);
/* ---- banded reduced system (csrc/ba_band.cu) -------------------------------------------------------------
 * S couples two free poses only if some patch is observed by both.  wv_ba_global_span reports the widest frame span of any
 * patch's observations among the free poses (after building the edge plan in `workspace` when build_plan != 0).  When
 * span <= bs <= wv_ba_band_max_bs() (bs >= 2), S is banded with a half-width of bs pose blocks:
 *   wv_ba_global_build_band accumulates it into a band buffer of wv_ba_band_floats(n6, bs) floats (about 6N * 6 (bs + 2) instead
 *   of (6N)^2 -- also what an all-reduce over ranks then moves); element (r, c), r >= c, lives at Sband[r * ld + c],
 *   ld = 6 (bs + 2) rounded up to a multiple of 4;
 *   wv_ba_band_solve factorises and solves it on the device (two launches, no host sync): damping S_dd <- S_dd (1 + 1e-4) + 1
 *   when damp != 0 (ba_cuda.cu:517-518), band Cholesky, dX = S^-1 y.  It replaces torch::linalg::cholesky + cholesky_solve on the
 *   dense matrix (ba_cuda.cu:521-523).  `factors`: wv_ba_band_factor_floats(n6, bs) floats of scratch, 16-byte aligned.
 *   A non-positive pivot sets WV_STATUS_NOT_SPD.  Everything else of an iteration (dz, retract) is unchanged. */
int wv_ba_band_max_bs(void);
size_t wv_ba_band_floats(int n6, int bs);
size_t wv_ba_band_factor_floats(int n6, int bs);
int wv_ba_global_span(const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses,
                      int64_t n_patches, int t0, int t1, int build_plan, int* span_out, void* workspace,
                      size_t workspace_bytes, int* status, wv_stream_t stream);
int wv_ba_global_build_band(float* poses, float* patches, const float* intrinsics, const float* target,
                            const float* weight, const float* lmbda, const int64_t* ii, const int64_t* jj,
                            const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0,
                            int t1, const float* patches_est, float mu, int shard, int nshards, int build_plan,
                            int bs, float* Sband, float* y, void* workspace, size_t workspace_bytes, int* status,
                            wv_stream_t stream);
int wv_ba_band_solve(const float* Sband, const float* y, int n6, int bs, int damp, float* factors, float* dX,
                     int* status, wv_stream_t stream);
/* S_dd <- S_dd * (1 + 1e-4) + 1 (ba_cuda.cu:517-518); n6 = 6 * (t1 - t0) */
int wv_ba_global_damp(float* S, int n6, wv_stream_t stream);
int wv_ba_global_dz(float* poses, float* patches, const float* intrinsics, const float* target,
                    const float* weight, const float* lmbda, const int64_t* ii, const int64_t* jj, int64_t E,
                    int64_t n_poses, int64_t n_patches, int P, int t0, int t1, const float* patches_est,
                    float mu, int shard, int nshards, const float* dX, float* dZ, void* workspace,
                    size_t workspace_bytes, int* status, wv_stream_t stream);
/* pose_retr_kernel + patch_retr_kernel (ba_cuda.cu:160-211) for poses t0..t1-1 and every patch of the plan */
int wv_ba_global_retract(float* poses, float* patches, int64_t E, int64_t n_poses, int64_t n_patches, int P,
                         int t0, int t1, const float* dX, const float* dZ, void* workspace,
                         size_t workspace_bytes, wv_stream_t stream);

/* cuda_ba.reproject (ba_cuda.cu:368-418, 543-575): coords [E][2][P][P], no depth clamp. */
int wv_ba_reproject(const float* poses, const float* patches, const float* intrinsics, const int64_t* ii,
                    const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches,
                    int P, float* coords, int* status, wv_stream_t stream);

/* ---- fused projective ops: what dpvo/projective_ops.py composes from lietorch group ops ----------
 * Layouts as the reference's tensors: poses [n_poses][7], patches [n_patches][3][P][P], intrinsics [n_poses][4] -- PER-FRAME
 * rows (intrinsics[ii] back-projects, intrinsics[jj] projects) and d = 1 / max(Z, 0.1), as projective_ops.py:19-49 does
 * (wv_ba_reproject is the fastba flavour: row 0 only, no clamp).
 * wv_pops_transform: transform(poses, patches, intrinsics, ii, jj, kk, depth, valid, jacobian=False, tonly)
 *   (projective_ops.py:53-104): coords [E][P][P][depth ? 3 : 2]; valid (nullable) [E][P][P] = (Z > 0.2).
 * wv_pops_flow_mag: flow_mag(..., beta) (projective_ops.py:111-121): out [E][P][P].
 * wv_pops_point_cloud: point_cloud(poses, patches, intrinsics, ix) (projective_ops.py:106-108): out [n][P][P][4].
 * wv_frame_distance: compute_keyframe_distance (dpvo/dpvo.py:383-408) for all pairs of the first n frames, patches of frame f
 *   at rows [f*M, (f+1)*M): D [n][n] symmetric, zero diagonal. */
int wv_pops_transform(const float* poses, const float* patches, const float* intrinsics, const int64_t* ii,
                      const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P,
                      int tonly, int depth, float* coords, float* valid, int* status, wv_stream_t stream);
/* the jacobian=True outputs of transform (projective_ops.py:71-99), centre pixel of every edge:
 *   Ji [E][2][6] = -Adj(Gij)^T Jj, Jj [E][2][6] = Jp Ja, Jz [E][2] = Jp Gij.matrix()[:, 3], validc [E] = (Z > 0.2);
 *   d = 1 / Z where |Z| > 0.2 else 0 (projective_ops.py:79-80), intrinsics of frame jj. */
int wv_pops_jacobian(const float* poses, const float* patches, const float* intrinsics, const int64_t* ii,
                     const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P,
                     int tonly, float* Ji, float* Jj, float* Jz, float* validc, int* status, wv_stream_t stream);
int wv_pops_flow_mag(const float* poses, const float* patches, const float* intrinsics, const int64_t* ii,
                     const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P,
                     float beta, float* out, int* status, wv_stream_t stream);
int wv_pops_point_cloud(const float* poses, const float* patches, const float* intrinsics, const int64_t* ix,
                        int64_t n, int64_t n_poses, int P, float* out, int* status, wv_stream_t stream);
int wv_frame_distance(const float* poses, const float* patches, const float* intrinsics, int64_t n, int64_t n_poses,
                      int64_t n_patches, int M, int P, float beta, float* D, wv_stream_t stream);

/* ---- SoftAgg of the update operator (dpvo/blocks.py:31-48, called from dpvo/net.py:88-89) ----------
 * wv_group_plan: `_, jx = torch.unique(ix, return_inverse=True)` (blocks.py:41) without the sort or the host sync.
 *   keys [E] int64 with key_min <= key < key_min + key_range (key_range <= 2^28; a key outside sets WV_STATUS_BAD_INDEX and
 *   belongs to no group) -> jx [E] int64 = rank of the key among the distinct keys present (torch.unique's sorted order),
 *   *n_groups (device word) = G, offsets [E + 1] int32 (entries 0..G valid: group g's members are order[offsets[g] ..
 *   offsets[g+1])), order [E] int32 = the edges grouped by jx, ascending edge index inside a group.
 * wv_soft_agg_forward: torch_scatter.scatter_softmax(gx, jx, dim=1) and scatter_sum(fx * w, jx, dim=1) (blocks.py:42-43;
 *   torch-scatter 2.1.2 composite/softmax.py: max per group, exp(src - max), sum, divide) in one pass:
 *   fx, gx [B][E][D] (dtype) -> y [B][g_capacity][D] (dtype), rows 0 .. min(G, g_capacity)-1 written.  fp32 arithmetic, members
 *   rows G .. g_capacity-1 zeroed (a caller that knows a bound on G -- patches in the buffer, frame pairs -- needs no host read
 *   of G: those rows are never gathered by jx).  fp32 arithmetic in a fixed order (deterministic; the reference's atomics are not). */
size_t wv_group_plan_workspace_bytes(int64_t E, int64_t key_range);
int wv_group_plan(const int64_t* keys, int64_t E, int64_t key_min, int64_t key_range, int64_t* jx, int32_t* order,
                  int32_t* offsets, int32_t* n_groups, void* workspace, size_t workspace_bytes, int* status,
                  wv_stream_t stream);
int wv_soft_agg_forward(const void* fx, const void* gx, const int32_t* order, const int32_t* offsets,
                        const int32_t* n_groups, void* y, int dtype, int B, int64_t E, int D, int64_t g_capacity,
                        wv_stream_t stream);

/* cuda_ba.neighbors (dpvo/fastba/ba.cpp:106-151), on the device and without host synchronisation:
 * for every edge the index of the previous / next edge of the same group `ii` in (jj, original index)
 * order, -1 at the ends.  ii values must lie in [0, n_groups). */
int wv_neighbors(const int64_t* ii, const int64_t* jj, int64_t E, int64_t n_groups, int64_t* ix,
                 int64_t* jx, void* workspace, size_t workspace_bytes, int* status, wv_stream_t stream);

/* wv_neighbors without a workspace argument, for callers that may not allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and results as wv_neighbors; the plan's scratch is a fixed 21 MiB static device array inside the library
 * (every GPU that loads it holds the array until the process exits), so the call allocates and frees nothing.
 *   limits: E <= WV_NEIGHBORS_STATIC_MAX_EDGES and n_groups <= WV_NEIGHBORS_STATIC_MAX_GROUPS, else WV_ELIMIT (nothing enqueued).
 *   all calls on one device share that array: calls that overlap on the GPU (different streams without an ordering between
 *   them) race -- serialise them.  The array keeps the previous call's contents; the call itself clears what must start at zero. */
#define WV_NEIGHBORS_STATIC_MAX_EDGES 262144
#define WV_NEIGHBORS_STATIC_MAX_GROUPS 1048576
int wv_neighbors_static(const int64_t* ii, const int64_t* jj, int64_t E, int64_t n_groups, int64_t* ix,
                        int64_t* jx, int* status, wv_stream_t stream);

/* cuda_ba.solve_system (dpvo/fastba/ba.cpp:156-234) -- HOST pointers, synchronous, fp64 like the
 * reference's Eigen::SimplicialCholesky<double>.  J_i / J_j [r][7][7], res [r][7] -> delta [n][7] with
 * n = max(ii, jj) + 1 (written to *n_out; pass delta == NULL to query n).  freen < 0 solves everything,
 * otherwise only the leading freen poses.  returns WV_OK, WV_EINVAL (i == j: the reference calls exit(1)),
 * or 1 when the damped normal matrix is not positive definite. */
int wv_solve_system_host(const float* J_i, const float* J_j, const int64_t* ii, const int64_t* jj,
                         const float* res, int64_t r, float ep, float lm, int freen, float* delta,
                         int64_t* n_out);

/* ---- COLMAP text export (dpvo/plot_utils.py:58-93) -- HOST pointers, synchronous, no device work ----------
 * The rows of points3D.txt / images.txt exactly as the reference's f-strings produce them (numbers in CPython's float repr),
 * written in one pass instead of Python `+=` per row.  Return the number of bytes written, WV_EWORKSPACE when `cap` is too
 * small (180 bytes per point row, 260 + name length per image row always suffice), or WV_EINVAL.
 *   wv_format_f64: repr(float(v)) into out32 (>= 32 bytes), returns its length -- the formatter itself, exported for tests.
 *   points: xyz [n][3] (the float32 point cloud widened to double, as .tolist() does), rgb [n][3] uint8, ids first_id ...
 *   images: pose7 [n][7] = (qw, qx, qy, qz, x, y, z) camera-from-world, names = n NUL-terminated strings back to back. */
int wv_format_f64(double v, char* out32);
int64_t wv_colmap_points3d_text(const double* xyz, const uint8_t* rgb, int64_t n, int64_t first_id, char* out, int64_t cap);
int64_t wv_colmap_images_text(const double* pose7, const char* names, int64_t n, char* out, int64_t cap);

/* ---------------------------------------------------------------------------------------------
 * host-buffer entry points (pinned or pageable HOST memory in, HOST memory out): the call a
 * non-torch embedder makes; copies are enqueued on `stream` around the same kernels and the call
 * returns after the stream has drained.  Used by bench.py's end-to-end leg.
 * ------------------------------------------------------------------------------------------- */
int wv_corr_forward_host(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii,
                         const int64_t* jj, void* out, int dtype, int B, int E, int C, int PH, int PW,
                         int n1, int n2, int H2, int W2, int radius, wv_stream_t stream);

int wv_ba_forward_host(float* poses, float* patches, const float* intrinsics, const float* target,
                       const float* weight, const float* lmbda, const int64_t* ii, const int64_t* jj,
                       const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0,
                       int t1, int iterations, const float* patches_est, float mu, int* status_out,
                       wv_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * host-buffer session: one tracker update() per call -- altcorr.corr on every pyramid level
 * (dpvo/dpvo.py:326-333: the same coords divided by 1, 4, ...), fastba.neighbors and fastba.BA on the
 * same edge list (dpvo/dpvo.py:711-749) -- with HOST buffers in and out.  The session keeps its
 * device buffers and three streams between calls; uploads, kernels and downloads of the different
 * pieces overlap (small inputs and the coarse level first, the big level-0 ring last, each result
 * downloaded while the next piece is still being uploaded).  Pinned host memory makes the copies
 * asynchronous; pageable memory works and serialises them.  The call returns when everything has
 * landed in the caller's buffers.
 * ------------------------------------------------------------------------------------------- */
typedef struct wv_session wv_session;

typedef struct {
  /* correlation */
  int dtype, B, E, C, P, radius, n1, n2, nlevels; /* nlevels <= 4 */
  const void* fmap1;                              /* [B][n1][C][P][P] */
  const void* fmap2[4];                           /* level l: [B][n2][C][H2[l]][W2[l]] */
  int H2[4], W2[4];
  float coord_scale[4];                           /* level l correlates at coords * coord_scale[l] (1, 0.25, ...) */
  const float* coords;                            /* [B][E][2][P][P] at level-0 resolution */
  const int64_t* ii1;                             /* [E] patch slot in fmap1, jj1: frame slot in fmap2 */
  const int64_t* jj1;
  void* corr_out[4];                              /* level l: [B][E][2r+1][2r+1][P][P] (dtype) */
  /* bundle adjustment + neighbors on the same E edges; iterations == 0 skips both */
  float* poses; float* patches;                   /* in/out */
  const float* intrinsics; const float* target; const float* weight; const float* lmbda;
  const int64_t* ii; const int64_t* jj; const int64_t* kk;
  int64_t n_poses, n_patches;
  int t0, t1, iterations;
  const float* patches_est; float mu;             /* nullable prior (see wv_ba_forward) */
  int64_t* nb_ix; int64_t* nb_jx;                 /* nullable: neighbors(kk, jj) */
  int status;                                     /* out: WV_STATUS_* bits */
  /* nullable, nlevels == 2, tracker configuration only: both levels as ONE volume [B][E][2r+1][2r+1][P][P][2] (DPVO.corr's
   * torch.stack layout, see wv_corr_forward_stacked) instead of corr_out[0..1]; WV_EUNSUPPORTED for other shapes */
  void* corr_stacked_out;
} wv_update_args;

wv_session* wv_session_create(void);              /* on the current device; NULL on failure */
void wv_session_destroy(wv_session* s);
int wv_session_update_host(wv_session* s, wv_update_args* a);

/* ---------------------------------------------------------------------------------------------
 * lietorch_backends -- the third module the tracker's host code imports by name (dpvo/lietorch/group_ops.py:1)
 * ------------------------------------------------------------------------------------------- */

/* group ids are the reference's (dpvo/lietorch/groups.py:236, 252, 268, 290) */
enum { WV_LIE_SO3 = 1, WV_LIE_RXSO3 = 2, WV_LIE_SE3 = 3, WV_LIE_SIM3 = 4 };
/* one id per m.def pair of dpvo/lietorch/src/lietorch.cpp:286-316 */
enum {
  WV_LIE_EXP = 0,       /* expm:  a [n][K]            -> X [n][N]                       lietorch_gpu.cu:20-43    */
  WV_LIE_LOG = 1,       /* logm:  X [n][N]            -> a [n][K]                                    :45-70    */
  WV_LIE_INV = 2,       /* inv:   X                   -> X^-1                                        :72-98    */
  WV_LIE_MUL = 3,       /* mul:   X, Y                -> X Y                                         :101-127  */
  WV_LIE_ADJ = 4,       /* adj:   X, a [n][K]         -> Adj(X) a                                    :129-158  */
  WV_LIE_ADJT = 5,      /* adjT:  X, a                -> Adj(X)^T a                                  :161-189  */
  WV_LIE_ACT = 6,       /* act:   X, p [n][3]         -> X p                                         :191-222  */
  WV_LIE_ACT4 = 7,      /* act4:  X, p [n][4]         -> X p (homogeneous)                           :225-257  */
  WV_LIE_MATRIX = 8,    /* as_matrix: X               -> [n][4][4] row-major                         :259-270  */
  WV_LIE_PROJECTOR = 9, /* projector: X               -> [n][N][N] row-major                         :272-281  */
  WV_LIE_JINV = 10      /* Jinv:  X, a                -> J_l^-1(Log X) a                             :283-294  */
};
/* Embedding N / tangent K per group: SO3 4/3, RxSO3 5/4, SE3 7/6, Sim3 8/7; storage (tx ty tz) (qx qy qz qw) (s).
 * dtype WV_F32 or WV_F64.  x_stride / y_stride: distance in SCALARS between consecutive batch items of the first / second
 * operand -- the dense value (N, K, 3 or 4) or 0 to broadcast one element over the batch (the reference materialises that with
 * repeat(), dpvo/lietorch/broadcasting.py:26-29).  on_host != 0: the pointers are HOST pointers and the call computes
 * synchronously on the calling thread (the reference's lietorch_cpu.cpp; same arithmetic, `stream` ignored).
 * wv_lie_backward follows the reference's gradient convention (row vectors over the first K entries of an N-wide row,
 * lietorch_gpu.cu:33-43); it writes dX [n][N] ([n][K] for expm) and, for binary ops, dY shaped like Y, all dense. */
int wv_lie_forward(int op, int group, int dtype, const void* X, int64_t x_stride, const void* Y, int64_t y_stride, void* out,
                   int64_t n, int on_host, wv_stream_t stream);
int wv_lie_backward(int op, int group, int dtype, const void* grad, const void* X, int64_t x_stride, const void* Y,
                    int64_t y_stride, void* dX, void* dY, int64_t n, int on_host, wv_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* WVB200_H_ */
