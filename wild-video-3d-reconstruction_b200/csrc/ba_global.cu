// Bundle adjustment with MANY free poses (the loop-closure / end-of-run global BA, dpvo/dpvo.py:436-505: t0 = 1, t1 = n,
// up to 2 047 free poses, ~1e7 edges) -- the case where the reference's cuda_ba materialises a dense 6N x Mp matrix E
// (19 GB at n = 2048, M = 192) and multiplies it by its transpose (1.2e14 flop), dpvo/fastba/ba_cuda.cu:443-444, 510-515.
//
// Here E never exists.  One warp per patch (segments of the edge plan, ba_plan.cuh), a lane per edge:
//   build   edge residuals / Jacobians -> the pose-block terms of B and v go straight into the dense reduced system
//           [S | y] (fp32 atomics on the LOWER triangle only; the source-frame block is warp-reduced first); the patch's
//           sparse column of E is a short list of (frame, 6-vector) entries in shared memory (one per observing frame,
//           <= ~26), C_k and u_k are warp sums; the patch is eliminated on the spot:
//           S[fa][fb] -= q_k E_a E_b^T over entry pairs, y[fa] -= q_k u_k E_a.   Patches can be sharded over ranks
//           (`shard` of `nshards`, contiguous runs of the plan's sorted unique patch ids): every rank then holds a PARTIAL
//           [S | y] and the caller sums them (one NCCL allreduce per Gauss-Newton iteration).
//   (caller) damping S_dd <- S_dd (1 + 1e-4) + 1 (ba_cuda.cu:517-518), dense Cholesky + solve (cuSOLVER through ATen in
//           the torch binding, exactly the reference's own two calls, :521-523).
//   dz      the same per-patch pass recomputes the entries and emits dZ_k = q_k (u_k - sum_a E_a . dX[fa]) (:524-525).
//   retract poses t0..t1-1 and all patches (:160-211).
// The accumulation order of the atomics is not fixed, so (like the reference, unlike the windowed kernel) results are
// reproducible only to rounding.
#include <algorithm>

#include "ba_edge.cuh"
#include "ba_plan.cuh"
#include "se3.cuh"

namespace wv {

static constexpr int kGThreads = 256;
static constexpr int kGWarps = kGThreads / 32;
static constexpr int kMaxEntries = 128;  // (frame, 6-vector) entries of one patch's column of E

struct GArgs {
  float* poses; float* patches; const float* intr; const float* target; const float* weight; const float* lmbda;
  const int64_t* ii; const int64_t* jj;
  const float* patches_est; float mu;
  int P, t0, N, n_poses;
  Plan p;
  float* S;         // [6N][6N] row-major, lower triangle (diagonal blocks: their lower triangle)
  float* y;         // [6N]
  const float* dX;  // [N][6]
  float* dZ;        // [Mp_cap], indexed by the plan's patch rank
  int shard, nshards;
  int* status;
};

// block (fa, fb), fa > fb: all 36 entries; fa == fb: lower triangle of the diagonal block.  val(r, c) = s * A[r] * B[c]
__device__ __forceinline__ void add_block(float* S, int n6, int fa, int fb, const float* A, const float* B, float s) {
  float* base = S + (size_t)(6 * fa) * n6 + 6 * fb;
  if (fa > fb) {
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
      for (int c = 0; c < 6; c++) atomicAdd(base + (size_t)r * n6 + c, s * A[r] * B[c]);
  } else {
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
      for (int c = 0; c <= r; c++) atomicAdd(base + (size_t)r * n6 + c, s * A[r] * B[c]);
  }
}

template <bool BUILD>
__global__ void __launch_bounds__(kGThreads) ba_global_kernel(GArgs a) {
  __shared__ float ent_v[kGWarps][kMaxEntries][6];
  __shared__ int ent_f[kGWarps][kMaxEntries];
  const int N = a.N, n6 = 6 * N;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int PP = a.P * a.P, ctr = (a.P / 2) * a.P + a.P / 2;
  const float fx = a.intr[0], fy = a.intr[1], cx = a.intr[2], cy = a.intr[3];  // row 0 only (ba_cuda.cu:232-238)
  const float lambda = a.lmbda[0];
  const int Mp = a.p.hdr->Mp;
  const int r_lo = (int)((long long)Mp * a.shard / a.nshards), r_hi = (int)((long long)Mp * (a.shard + 1) / a.nshards);
  const int gw = blockIdx.x * kGWarps + warp, nw = gridDim.x * kGWarps;
  const unsigned FULL = 0xffffffffu;
  float(*ev)[6] = ent_v[warp];
  int* ef = ent_f[warp];

  for (int r = r_lo + gw; r < r_hi; r += nw) {
    const int id = a.p.kx[r], off = a.p.seg_off[r], n = a.p.seg_off[r + 1] - off;
    const float* pc = a.patches + (size_t)id * 3 * PP;
    const float px = pc[ctr], py = pc[PP + ctr], pd = pc[2 * PP + ctr];
    float Csum = 0.f, usum = 0.f;
    int L = 0;  // entries so far (warp-uniform)
    bool overflow = false;
    for (int base = 0; base < n; base += 32) {
      bool act = base + lane < n;
      const int e = act ? a.p.eperm[off + base + lane] : 0;
      int ix = -1, jx = -1;
      if (act) {
        const int64_t iv = a.ii[e], jv = a.jj[e];
        if (iv < 0 || iv >= a.n_poses || jv < 0 || jv >= a.n_poses) { set_status(a.status, WV_STATUS_BAD_INDEX); act = false; }
        else { ix = (int)iv; jx = (int)jv; }
      }
      const int fi = ix - a.t0, fj = jx - a.t0;
      const bool free_i = act && fi >= 0 && fi < N, free_j = act && fj >= 0 && fj < N;
      EdgeTerms T;
      float Ej[6], Ei[6];
#pragma unroll
      for (int c = 0; c < 6; c++) { Ej[c] = 0.f; Ei[c] = 0.f; }
      if (act) {
        edge_terms(a, e, ix, jx, px, py, pd, fx, fy, cx, cy, T);
        Csum += T.w[0] * T.Jz[0] * T.Jz[0] + T.w[1] * T.Jz[1] * T.Jz[1];
        usum += T.w[0] * T.r[0] * T.Jz[0] + T.w[1] * T.r[1] * T.Jz[1];
#pragma unroll
        for (int c = 0; c < 6; c++) {
          Ej[c] = T.w[0] * T.Jz[0] * T.Jj[0][c] + T.w[1] * T.Jz[1] * T.Jj[1][c];
          Ei[c] = -(T.w[0] * T.Jz[0] * T.Ji[0][c] + T.w[1] * T.Jz[1] * T.Ji[1][c]);
        }
      }
      // source frame shared by the whole chunk (the tracker's edge lists): its block / entry are reduced in the warp first
      const int ix0 = __shfl_sync(FULL, ix, 0);
      const bool uni = __all_sync(FULL, !act || ix == ix0);
      const int f0 = ix0 - a.t0;
      const bool free0 = uni && f0 >= 0 && f0 < N;
      if (BUILD) {
        if (act) {
          if (free_j) {  // (j, j) block and v_j: both residual rows combined, one set of atomics per edge
            float* base_jj = a.S + (size_t)(6 * fj) * n6 + 6 * fj;
#pragma unroll
            for (int p = 0; p < 6; p++) {
#pragma unroll
              for (int c = 0; c <= p; c++)
                atomicAdd(base_jj + (size_t)p * n6 + c, T.w[0] * T.Jj[0][p] * T.Jj[0][c] + T.w[1] * T.Jj[1][p] * T.Jj[1][c]);
              atomicAdd(a.y + 6 * fj + p, T.w[0] * T.r[0] * T.Jj[0][p] + T.w[1] * T.r[1] * T.Jj[1][p]);
            }
          }
          if (free_i && free_j) {  // (i, j) cross block: -w Ji Jj^T
            if (fi > fj) {
              float* b2 = a.S + (size_t)(6 * fi) * n6 + 6 * fj;
#pragma unroll
              for (int p = 0; p < 6; p++)
#pragma unroll
                for (int c = 0; c < 6; c++)
                  atomicAdd(b2 + (size_t)p * n6 + c, -(T.w[0] * T.Ji[0][p] * T.Jj[0][c] + T.w[1] * T.Ji[1][p] * T.Jj[1][c]));
            } else if (fi < fj) {
              float* b2 = a.S + (size_t)(6 * fj) * n6 + 6 * fi;
#pragma unroll
              for (int p = 0; p < 6; p++)
#pragma unroll
                for (int c = 0; c < 6; c++)
                  atomicAdd(b2 + (size_t)c * n6 + p, -(T.w[0] * T.Ji[0][p] * T.Jj[0][c] + T.w[1] * T.Ji[1][p] * T.Jj[1][c]));
            } else {
              float* b2 = a.S + (size_t)(6 * fi) * n6 + 6 * fi;
#pragma unroll
              for (int p = 0; p < 6; p++)
#pragma unroll
                for (int c = 0; c <= p; c++)
                  atomicAdd(b2 + (size_t)p * n6 + c, -(T.w[0] * (T.Ji[0][p] * T.Jj[0][c] + T.Jj[0][p] * T.Ji[0][c]) +
                                                       T.w[1] * (T.Ji[1][p] * T.Jj[1][c] + T.Jj[1][p] * T.Ji[1][c])));
            }
          }
        }
        // (i, i) block and v_i
        if (free0) {
          int t = 0;
#pragma unroll
          for (int p = 0; p < 6; p++) {
#pragma unroll
            for (int c = 0; c <= p; c++) {
              float v = act ? T.w[0] * T.Ji[0][p] * T.Ji[0][c] + T.w[1] * T.Ji[1][p] * T.Ji[1][c] : 0.f;
              v = warp_sum(v);
              if (lane == t) atomicAdd(a.S + (size_t)(6 * f0 + p) * n6 + 6 * f0 + c, v);
              t++;
            }
            float w = act ? T.w[0] * T.r[0] * T.Ji[0][p] + T.w[1] * T.r[1] * T.Ji[1][p] : 0.f;
            w = warp_sum(w);
            if (lane == 21 + p) atomicAdd(a.y + 6 * f0 + p, -w);
          }
        } else if (free_i) {  // mixed source frames: per-edge atomics
          float* b2 = a.S + (size_t)(6 * fi) * n6 + 6 * fi;
#pragma unroll
          for (int p = 0; p < 6; p++) {
#pragma unroll
            for (int c = 0; c <= p; c++)
              atomicAdd(b2 + (size_t)p * n6 + c, T.w[0] * T.Ji[0][p] * T.Ji[0][c] + T.w[1] * T.Ji[1][p] * T.Ji[1][c]);
            atomicAdd(a.y + 6 * fi + p, -(T.w[0] * T.r[0] * T.Ji[0][p] + T.w[1] * T.r[1] * T.Ji[1][p]));
          }
        }
      }
      // ---- entries of the patch's column of E
      {
        const unsigned mj = __ballot_sync(FULL, free_j);
        const int nj = __popc(mj);
        if (L + nj <= kMaxEntries) {
          if (free_j) {
            const int s = L + __popc(mj & ((1u << lane) - 1u));
            ef[s] = fj;
#pragma unroll
            for (int c = 0; c < 6; c++) ev[s][c] = Ej[c];
          }
          L += nj;
        } else overflow = true;
        if (uni) {
          float Es[6];
#pragma unroll
          for (int c = 0; c < 6; c++) Es[c] = warp_sum(Ei[c]);
          if (free0) {
            if (L + 1 <= kMaxEntries) {
              if (lane == 0) {
                ef[L] = f0;
#pragma unroll
                for (int c = 0; c < 6; c++) ev[L][c] = Es[c];
              }
              L += 1;
            } else overflow = true;
          }
        } else {
          const unsigned mi = __ballot_sync(FULL, free_i);
          const int ni = __popc(mi);
          if (L + ni <= kMaxEntries) {
            if (free_i) {
              const int s = L + __popc(mi & ((1u << lane) - 1u));
              ef[s] = fi;
#pragma unroll
              for (int c = 0; c < 6; c++) ev[s][c] = Ei[c];
            }
            L += ni;
          } else overflow = true;
        }
      }
      __syncwarp();
    }
    if (overflow && lane == 0) set_status(a.status, WV_STATUS_TOO_DENSE);
    float C = warp_sum(Csum), u = warp_sum(usum);
    if (a.patches_est) {  // depth prior (dpvo/ba.py:151-159)
      const float dest = a.patches_est[(size_t)id * 3 * PP + 2 * PP + ctr];
      if (dest > 0.f) { C += a.mu; u -= a.mu * (pd - dest); }
    }
    const float q = 1.f / (C + lambda);
    if (BUILD) {
      if (N > 0) {
        for (int t = lane; t < L * L; t += 32) {
          const int ea = t / L, eb = t - ea * L;
          const int fa = ef[ea], fb = ef[eb];
          if (fa >= fb) add_block(a.S, n6, fa, fb, ev[ea], ev[eb], -q);
        }
        for (int t = lane; t < L * 6; t += 32) atomicAdd(a.y + 6 * ef[t / 6] + t % 6, -q * u * ev[t / 6][t % 6]);
      }
    } else {
      float sdot = 0.f;
      for (int t = lane; t < L * 6; t += 32) sdot += ev[t / 6][t % 6] * a.dX[6 * ef[t / 6] + t % 6];
      sdot = warp_sum(sdot);
      if (lane == 0) a.dZ[r] = q * (u - sdot);
    }
    __syncwarp();
  }
}

__global__ void ba_global_damp_kernel(float* S, int n6) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d < n6) { float* p = S + (size_t)d * n6 + d; const float s = *p; *p = s + (s * 1e-4f + 1.0f); }  // ba_cuda.cu:517-518
}

__global__ void ba_global_retract_kernel(GArgs a) {
  const int PP = a.P * a.P;
  const int Mp = a.p.hdr->Mp;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < Mp + a.N; t += gridDim.x * blockDim.x) {
    if (t < Mp) {  // patch_retr_kernel, ba_cuda.cu:191-211
      float* pdz = a.patches + (size_t)a.p.kx[t] * 3 * PP + 2 * PP;
      float dd = pdz[0] + a.dZ[t];
      dd = (dd > 20.f) ? 1.0f : dd;
      dd = fmaxf(dd, 1e-4f);
      for (int k = 0; k < PP; k++) pdz[k] = dd;
    } else {       // pose_retr_kernel, ba_cuda.cu:160-188
      const int f = t - Mp;
      float* Pp = a.poses + 7 * (size_t)(a.t0 + f);
      Vec3 tt{Pp[0], Pp[1], Pp[2]};
      Quat q{Pp[3], Pp[4], Pp[5], Pp[6]};
      se3_retract(a.dX + 6 * f, tt, q);
      Pp[0] = tt.x; Pp[1] = tt.y; Pp[2] = tt.z; Pp[3] = q.x; Pp[4] = q.y; Pp[5] = q.z; Pp[6] = q.w;
    }
  }
}

// ---------------------------------------------------------------------------------------------------
int plan_grid();  // ba_plan.cu

size_t ba_global_workspace_bytes(int64_t E, int64_t n_patches) { return plan_bytes(E, n_patches, plan_grid()) + 1024; }

static GArgs make_gargs(float* poses, float* patches, const float* intr, const float* target, const float* weight, const float* lmbda,
                        const int64_t* ii, const int64_t* jj, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0, int t1,
                        const float* patches_est, float mu, void* ws, int shard, int nshards, int* status) {
  GArgs a{};
  plan_carve(&a.p, ws, E, n_patches, plan_grid());
  a.poses = poses; a.patches = patches; a.intr = intr; a.target = target; a.weight = weight; a.lmbda = lmbda; a.ii = ii; a.jj = jj;
  a.patches_est = patches_est; a.mu = mu; a.P = P; a.t0 = t0; a.N = t1 - t0; a.n_poses = (int)n_poses;
  a.shard = shard; a.nshards = nshards; a.status = status;
  return a;
}

static int global_grid(int64_t n_patches_guess) {
  const int64_t want = (n_patches_guess + kGWarps - 1) / kGWarps;
  const int64_t cap = (int64_t)sm_count() * 8;
  return (int)std::max<int64_t>(1, std::min(want, cap));
}

int ba_global_build(float* poses, float* patches, const float* intr, const float* target, const float* weight, const float* lmbda,
                    const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P,
                    int t0, int t1, const float* patches_est, float mu, int shard, int nshards, int build_plan, float* S, float* y,
                    void* ws, size_t ws_bytes, int* status, cudaStream_t st) {
  if (ws_bytes < ba_global_workspace_bytes(E, n_patches)) return WV_EWORKSPACE;
  GArgs a = make_gargs(poses, patches, intr, target, weight, lmbda, ii, jj, E, n_poses, n_patches, P, t0, t1, patches_est, mu, ws,
                       shard, nshards, status);
  const size_t n6 = 6 * (size_t)a.N;
  a.S = S; a.y = y;
  if (build_plan) {
    const int rc = plan_build(a.p, kk, jj, ii, n_poses, nullptr, nullptr, status, st);
    if (rc != WV_OK) return rc;
  }
  if (n6) {
    WV_CUDA_TRY(cudaMemsetAsync(S, 0, n6 * n6 * sizeof(float), st));
    WV_CUDA_TRY(cudaMemsetAsync(y, 0, n6 * sizeof(float), st));
  }
  ba_global_kernel<true><<<global_grid(std::min(E, n_patches)), kGThreads, 0, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int ba_global_damp(float* S, int n6, cudaStream_t st) {
  if (n6 <= 0) return WV_OK;
  ba_global_damp_kernel<<<(n6 + 255) / 256, 256, 0, st>>>(S, n6);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int ba_global_dz(float* poses, float* patches, const float* intr, const float* target, const float* weight, const float* lmbda,
                 const int64_t* ii, const int64_t* jj, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0, int t1,
                 const float* patches_est, float mu, int shard, int nshards, const float* dX, float* dZ, void* ws, size_t ws_bytes,
                 int* status, cudaStream_t st) {
  if (ws_bytes < ba_global_workspace_bytes(E, n_patches)) return WV_EWORKSPACE;
  GArgs a = make_gargs(poses, patches, intr, target, weight, lmbda, ii, jj, E, n_poses, n_patches, P, t0, t1, patches_est, mu, ws,
                       shard, nshards, status);
  a.dX = dX; a.dZ = dZ;
  WV_CUDA_TRY(cudaMemsetAsync(dZ, 0, sizeof(float) * (size_t)std::min(E, n_patches), st));
  ba_global_kernel<false><<<global_grid(std::min(E, n_patches)), kGThreads, 0, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int ba_global_retract(float* poses, float* patches, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0, int t1,
                      const float* dX, const float* dZ, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (ws_bytes < ba_global_workspace_bytes(E, n_patches)) return WV_EWORKSPACE;
  GArgs a = make_gargs(poses, patches, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, E, n_poses, n_patches, P, t0, t1, nullptr,
                       0.f, ws, 0, 1, nullptr);
  a.dX = dX; a.dZ = const_cast<float*>(dZ);
  ba_global_retract_kernel<<<global_grid(std::min(E, n_patches) / 32 + 1), kGThreads, 0, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

}  // namespace wv
