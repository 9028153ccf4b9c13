// Sliding-window bundle adjustment: ONE persistent cooperative kernel per cuda_ba.forward call.
//
// Replaces, per Gauss-Newton iteration, the reference's 5 zero_() launches, the atomics edge kernel
// (dpvo/fastba/ba_cuda.cu:214-365: up to 340 global fp32 atomicAdds per edge onto a 6N x 6N matrix), the
// ~20 ATen launches of the Schur complement / Cholesky (:492-526) and both retraction kernels (:160-211).
//
// Design (B200: 148 SMs, 1 CTA each, everything between HBM reads lives in shared memory):
//   phase A  one warp per patch segment of the plan (ba_plan.cuh); a lane owns one edge.  The patch-level sums
//            C_k, u_k, E_k are warp reductions; pose-block contributions go to a PER-WARP lower-triangular
//            table in shared memory (lanes of a patch hit distinct target blocks -> plain read-modify-write,
//            no atomics); the patch is eliminated on the spot, S_w -= q_k E_k E_k^T, so the dense 6N x Mp
//            matrix E of the reference is never formed.  Warp tables are summed in fixed order per CTA.
//   phase B  CTA partials are summed in fixed CTA order (grid-strided over the 6N(6N+1)/2 + 6N entries).
//   phase C  every CTA factorises the same damped S in shared memory (redundant Cholesky: no broadcast, no
//            extra grid sync, bitwise identical everywhere), back-substitutes its own patches' depths and
//            retracts them; CTA 0 retracts the poses.
// Summation order is a pure function of the edge list, so results are bitwise reproducible run to run
// (the reference's atomics are not).  fp32 throughout, like the reference.
#include <cooperative_groups.h>
#include <stdlib.h>

#include <algorithm>

#include "ba_edge.cuh"
#include "ba_plan.cuh"
#include "se3.cuh"

namespace cg = cooperative_groups;

namespace wv {

static constexpr int kBaThreads = 256;
static constexpr int kBaWarps = kBaThreads / 32;
static constexpr int kMaxFree = 16;  // 6N <= 96: per-warp tables must fit shared memory

struct BAArgs {
  float* poses; float* patches; const float* intr; const float* target; const float* weight; const float* lmbda;
  const int64_t* ii; const int64_t* jj;
  const float* patches_est; float mu;
  int P, t0, N, iterations;
  int n_poses, n_patches;
  const int64_t* kk_check;  // non-null when the plan was built by an earlier call (wv_neighbors): every edge is verified against it
  Plan p;
  float* partial;  // [grid][nS]
  float* Sfull;    // [nS]
  float* Ecols;    // [Mp_cap][6N]
  float* qk;       // [Mp_cap]
  float* uk;       // [Mp_cap]
  int* status;
};

__device__ __forceinline__ int tri(int r, int c) { return ((r * (r + 1)) >> 1) + c; }  // r >= c

// One residual row of one edge: add its pose-block terms to the warp table.
//   a = (-Ji at block i, +Jj at block j); B += w a a^T, v += w r a   (ba_cuda.cu:293-316)
// with_i: also add the (i,i) block / v_i here (serialised slow path); otherwise the caller reduces them.
__device__ __forceinline__ void scatter_row(float* S, float* y, int fi, int fj, bool free_i, bool free_j,
                                            const float* Ji, const float* Jj, float w, float r, bool with_i) {
  if (free_j) {
    const int b = 6 * fj;
#pragma unroll
    for (int a = 0; a < 6; a++) {
#pragma unroll
      for (int c = 0; c <= a; c++) S[tri(b + a, b + c)] += w * Jj[a] * Jj[c];
      y[b + a] += w * r * Jj[a];
    }
  }
  if (free_i && free_j) {
    if (fi == fj) {
      const int b = 6 * fi;
#pragma unroll
      for (int a = 0; a < 6; a++)
#pragma unroll
        for (int c = 0; c <= a; c++) S[tri(b + a, b + c)] -= w * (Ji[a] * Jj[c] + Jj[a] * Ji[c]);
    } else if (fi > fj) {
#pragma unroll
      for (int a = 0; a < 6; a++)
#pragma unroll
        for (int c = 0; c < 6; c++) S[tri(6 * fi + a, 6 * fj + c)] -= w * Ji[a] * Jj[c];
    } else {
#pragma unroll
      for (int a = 0; a < 6; a++)
#pragma unroll
        for (int c = 0; c < 6; c++) S[tri(6 * fj + c, 6 * fi + a)] -= w * Ji[a] * Jj[c];
    }
  }
  if (with_i && free_i) {
    const int b = 6 * fi;
#pragma unroll
    for (int a = 0; a < 6; a++) {
#pragma unroll
      for (int c = 0; c <= a; c++) S[tri(b + a, b + c)] += w * Ji[a] * Ji[c];
      y[b + a] -= w * r * Ji[a];
    }
  }
}

__global__ void __launch_bounds__(kBaThreads, 1) ba_window_kernel(BAArgs a) {
  extern __shared__ float sm[];
  cg::grid_group grid = cg::this_grid();
  const int N = a.N, n6 = 6 * N, nT = (n6 * (n6 + 1)) >> 1, nS = nT + n6;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* tabs = sm;                      // [kBaWarps][nS]  (S lower triangle | y)
  float* myS = tabs + warp * nS;
  float* myY = myS + nT;
  float* myE = sm + kBaWarps * nS + warp * n6;  // E_k of the patch in flight
  const int PP = a.P * a.P, ctr = (a.P / 2) * a.P + a.P / 2;
  const float fx = a.intr[0], fy = a.intr[1], cx = a.intr[2], cy = a.intr[3];  // row 0 only (ba_cuda.cu:232-238)
  const float lambda = a.lmbda[0];
  const int Mp = a.p.hdr->Mp;
  const int per = (Mp + (int)gridDim.x - 1) / (int)gridDim.x;
  const int r_lo = min(Mp, (int)blockIdx.x * per), r_hi = min(Mp, r_lo + per);
  const unsigned FULL = 0xffffffffu;

  for (int itr = 0; itr < a.iterations; itr++) {
    for (int t = threadIdx.x; t < kBaWarps * nS; t += blockDim.x) tabs[t] = 0.f;
    __syncthreads();

    // ------------------------------------------------------------------ phase A: edges -> tables
    for (int r = r_lo + warp; r < r_hi; r += kBaWarps) {
      int id = a.p.kx[r];
      const int off = a.p.seg_off[r];
      int n = a.p.seg_off[r + 1] - off;
      if (id >= a.n_patches) {  // only possible on a reused plan (its ids were checked against the plan's group count)
        if (lane == 0) set_status(a.status, WV_STATUS_BAD_INDEX);
        id = 0; n = 0;
      }
      const float* pc = a.patches + (size_t)id * 3 * PP;
      const float px = pc[ctr], py = pc[PP + ctr], pd = pc[2 * PP + ctr];
      for (int c = lane; c < n6; c += 32) myE[c] = 0.f;
      __syncwarp();
      float Csum = 0.f, usum = 0.f;
      for (int base = 0; base < n; base += 32) {
        bool act = base + lane < n;
        const int e = act ? a.p.eperm[off + base + lane] : 0;
        int ix = -1, jx = -2 - lane;
        if (act) {  // a plan built by neighbors() has not looked at ii / jj: range-check them here
          const int64_t iv = a.ii[e], jv = a.jj[e];
          if (iv < 0 || iv >= a.n_poses || jv < 0 || jv >= a.n_poses) { set_status(a.status, WV_STATUS_BAD_INDEX); act = false; }
          else if (a.kk_check && a.kk_check[e] != (int64_t)id) { set_status(a.status, WV_STATUS_STALE_PLAN); act = false; }
          else { ix = (int)iv; jx = (int)jv; }
        }
        const int fi = ix - a.t0, fj = jx - a.t0;
        const bool free_i = act && fi >= 0 && fi < N, free_j = act && fj >= 0 && fj < N;
        EdgeTerms T;
        if (act) {
          edge_terms(a, e, ix, jx, px, py, pd, fx, fy, cx, cy, T);
          Csum += T.w[0] * T.Jz[0] * T.Jz[0] + T.w[1] * T.Jz[1] * T.Jz[1];
          usum += T.w[0] * T.r[0] * T.Jz[0] + T.w[1] * T.r[1] * T.Jz[1];
        }
        if (N == 0) continue;
        // fast path: one source frame for the whole chunk, all target frames distinct
        const int ix0 = __shfl_sync(FULL, ix, 0);
        const bool uni = __all_sync(FULL, !act || ix == ix0);
        const unsigned same = __match_any_sync(FULL, jx);
        const bool dup = __any_sync(FULL, act && (same & ~(1u << lane)) != 0u);
        if (uni && !dup) {
          float II[21], vi[6], Ei[6];
#pragma unroll
          for (int t = 0; t < 21; t++) II[t] = 0.f;
#pragma unroll
          for (int t = 0; t < 6; t++) { vi[t] = 0.f; Ei[t] = 0.f; }
          if (act) {
#pragma unroll
            for (int row = 0; row < 2; row++) {
              scatter_row(myS, myY, fi, fj, free_i, free_j, T.Ji[row], T.Jj[row], T.w[row], T.r[row], false);
              if (free_j) {
#pragma unroll
                for (int c = 0; c < 6; c++) myE[6 * fj + c] += T.w[row] * T.Jz[row] * T.Jj[row][c];
              }
              if (free_i) {
                int t = 0;
#pragma unroll
                for (int p = 0; p < 6; p++) {
#pragma unroll
                  for (int c = 0; c <= p; c++) II[t++] += T.w[row] * T.Ji[row][p] * T.Ji[row][c];
                  vi[p] += T.w[row] * T.r[row] * T.Ji[row][p];
                  Ei[p] += T.w[row] * T.Jz[row] * T.Ji[row][p];
                }
              }
            }
          }
          __syncwarp();
          const int f0 = ix0 - a.t0;
          if (f0 >= 0 && f0 < N) {  // (i,i) block, v_i, E_i: warp-reduced, then added by one lane each
            const int b = 6 * f0;
            int t = 0;
#pragma unroll
            for (int p = 0; p < 6; p++)
#pragma unroll
              for (int c = 0; c <= p; c++) {
                const float v = warp_sum(II[t]);
                if (lane == t) myS[tri(b + p, b + c)] += v;
                t++;
              }
#pragma unroll
            for (int p = 0; p < 6; p++) {
              const float v = warp_sum(vi[p]);
              if (lane == 21 + p) myY[b + p] -= v;
              const float w = warp_sum(Ei[p]);
              if (lane == p) myE[b + p] -= w;
            }
          }
          __syncwarp();
        } else {  // general graph: serialise the lanes of this chunk (deterministic, rare)
          for (int l = 0; l < 32; l++) {
            if (lane == l && act) {
#pragma unroll
              for (int row = 0; row < 2; row++) {
                scatter_row(myS, myY, fi, fj, free_i, free_j, T.Ji[row], T.Jj[row], T.w[row], T.r[row], true);
#pragma unroll
                for (int c = 0; c < 6; c++) {
                  if (free_j) myE[6 * fj + c] += T.w[row] * T.Jz[row] * T.Jj[row][c];
                  if (free_i) myE[6 * fi + c] -= T.w[row] * T.Jz[row] * T.Ji[row][c];
                }
              }
            }
            __syncwarp();
          }
        }
      }
      // ---- eliminate the patch: C_k, u_k (+ depth prior, dpvo/ba.py:151-159), q_k, Schur update
      float C = warp_sum(Csum), u = warp_sum(usum);
      if (a.patches_est) {
        const float dest = a.patches_est[(size_t)id * 3 * PP + 2 * PP + ctr];
        if (dest > 0.f) { C += a.mu; u -= a.mu * (pd - dest); }
      }
      const float q = 1.f / (C + lambda);
      if (lane == 0) { a.qk[r] = q; a.uk[r] = u; }
      if (N > 0) {
        for (int c = lane; c < n6; c += 32) a.Ecols[(size_t)r * n6 + c] = myE[c];
        for (int pr = lane; pr < (n6 >> 1); pr += 32) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int rr = h ? n6 - 1 - pr : pr;
            const float er = q * myE[rr];
            float* Sr = myS + tri(rr, 0);
            for (int c = 0; c <= rr; c++) Sr[c] -= er * myE[c];
            myY[rr] -= er * u;
          }
        }
      }
      __syncwarp();
    }
    __syncthreads();

    bool failed = false;
    if (N > 0) {
      // ---------------------------------------------------------------- phase B: fixed-order reductions
      for (int t = threadIdx.x; t < nS; t += blockDim.x) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kBaWarps; w++) s += tabs[w * nS + t];
        a.partial[(size_t)blockIdx.x * nS + t] = s;
      }
      grid.sync();
      {
        // cross-CTA sum in a fixed order: 16 lanes per element, lane s adds partials s, s+16, ... in order, then a fixed
        // xor-shuffle tree over the 16 lanes (bitwise reproducible; 37 888 threads share the 1 890 x 148 loads)
        const int g = blockIdx.x * blockDim.x + threadIdx.x, grp = g >> 4, sub = g & 15;
        const int ngrp = (gridDim.x * blockDim.x) >> 4;
        for (int base = 0; base < nS; base += ngrp) {
          const int t = base + grp;
          float s = 0.f;
          if (t < nS)
            for (int c = sub; c < (int)gridDim.x; c += 16) s += a.partial[(size_t)c * nS + t];
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
          if (t < nS && sub == 0) a.Sfull[t] = s;
        }
      }
      grid.sync();
      // a reused plan that does not belong to this kk (flagged by some CTA in phase A, visible to all after the grid syncs): stop
      // BEFORE anything is applied -- poses and patches are still the caller's, who rebuilds the plan and calls again
      if (a.kk_check && (*reinterpret_cast<volatile int*>(a.status) & WV_STATUS_STALE_PLAN)) break;
      // ---------------------------------------------------------------- phase C: damped Cholesky (every CTA)
      // Blocked right-looking LL^T on the (n6+1) x n6 trapezoid [S ; y^T] in shared memory, one 6-column pose block per step
      // (N steps of 2 barriers instead of 6N column steps): carrying the right-hand side as an extra row makes the forward
      // substitution L z = y fall out of the factorisation.  L overwrites the lower triangle; dg holds 1 / L_cc.
      const int ld = n6 + 1;           // odd: column walks are bank-conflict free
      float* A = sm;                   // [n6 + 1][ld]
      float* dg = A + (n6 + 1) * ld;   // [n6]  reciprocal diagonal of L
      float* bx = dg + n6;             // [n6]  the solution dX
      for (int t = threadIdx.x; t < nT; t += blockDim.x) {  // packed lower triangle, coalesced
        int rr = (int)((sqrtf(8.f * (float)t + 1.f) - 1.f) * 0.5f);
        while (tri(rr + 1, 0) <= t) rr++;
        while (tri(rr, 0) > t) rr--;
        const int cc = t - tri(rr, 0);
        float s = a.Sfull[t];
        if (rr == cc) s = s + (s * 1e-4f + 1.0f);  // S += I * (1e-4 * S + 1.0), ba_cuda.cu:517-518
        A[rr * ld + cc] = s;
      }
      for (int t = threadIdx.x; t < n6; t += blockDim.x) A[n6 * ld + t] = a.Sfull[nT + t];
      __syncthreads();
      const int ti = threadIdx.x >> 4, tk = threadIdx.x & 15;  // 16 x 16 threads over (row, column) of the trailing update
      for (int J = 0; J < N; J++) {
        const int c0 = 6 * J;
        // (a) the 6x6 diagonal block: EVERY thread factorises it in its own registers (same inputs, same arithmetic), so the
        // panel solve below needs no barrier after it; thread 0 files the result once the panel no longer reads the original
        float Lb[6][6], inv[6];
#pragma unroll
        for (int r = 0; r < 6; r++)
#pragma unroll
          for (int c = 0; c <= r; c++) Lb[r][c] = A[(c0 + r) * ld + c0 + c];
        bool bad = false;
#pragma unroll
        for (int c = 0; c < 6; c++) {
          float d = Lb[c][c];
#pragma unroll
          for (int m = 0; m < c; m++) d = fmaf(-Lb[c][m], Lb[c][m], d);
          if (!(d > 0.f)) bad = true;
          const float dj = sqrtf(d);
          inv[c] = 1.f / dj;
          Lb[c][c] = dj;
#pragma unroll
          for (int r = c + 1; r < 6; r++) {
            float v = Lb[r][c];
#pragma unroll
            for (int m = 0; m < c; m++) v = fmaf(-Lb[r][m], Lb[c][m], v);
            Lb[r][c] = v * inv[c];
          }
        }
        if (bad) { failed = true; break; }  // same value in every thread of every CTA
        // (b) panel: rows below the block (and the rhs row) x L_JJ^-T, one thread per row
        for (int i = c0 + 6 + threadIdx.x; i <= n6; i += blockDim.x) {
          float x[6];
#pragma unroll
          for (int c = 0; c < 6; c++) x[c] = A[i * ld + c0 + c];
#pragma unroll
          for (int c = 0; c < 6; c++) {
#pragma unroll
            for (int m = 0; m < c; m++) x[c] = fmaf(-x[m], Lb[c][m], x[c]);
            x[c] *= inv[c];
          }
#pragma unroll
          for (int c = 0; c < 6; c++) A[i * ld + c0 + c] = x[c];
        }
        __syncthreads();
        if (threadIdx.x == 0) {
#pragma unroll
          for (int r = 0; r < 6; r++) {
#pragma unroll
            for (int c = 0; c <= r; c++) A[(c0 + r) * ld + c0 + c] = Lb[r][c];
            dg[c0 + r] = inv[r];
          }
        }
        // (c) trailing update A_ik -= L_i,J . L_k,J for k > block, i >= k (i runs over the rhs row too)
        const int first = c0 + 6;
        for (int i = first + ti; i <= n6; i += 16) {
          float li[6];
#pragma unroll
          for (int c = 0; c < 6; c++) li[c] = A[i * ld + c0 + c];
          const int kend = i < n6 ? i : n6 - 1;
          for (int k = first + tk; k <= kend; k += 16) {
            float acc = A[i * ld + k];
#pragma unroll
            for (int c = 0; c < 6; c++) acc = fmaf(-li[c], A[k * ld + c0 + c], acc);
            A[i * ld + k] = acc;
          }
        }
        __syncthreads();
      }
      if (failed) {
        if (blockIdx.x == 0 && threadIdx.x == 0) set_status(a.status, WV_STATUS_NOT_SPD);
        break;
      }
      if (warp == 0) {  // L^T x = z: z (row n6) lives in registers, lane l owns entries l, l+32, l+64; row i of L is read ahead
        float z0 = lane < n6 ? A[n6 * ld + lane] : 0.f, z1 = lane + 32 < n6 ? A[n6 * ld + lane + 32] : 0.f,
              z2 = lane + 64 < n6 ? A[n6 * ld + lane + 64] : 0.f;
#pragma unroll 4
        for (int i = n6 - 1; i >= 0; i--) {
          const float* Li = A + i * ld;
          const float l0 = lane < i ? Li[lane] : 0.f, l1 = lane + 32 < i ? Li[lane + 32] : 0.f, l2 = lane + 64 < i ? Li[lane + 64] : 0.f;
          const float zi = __shfl_sync(0xffffffffu, i < 32 ? z0 : (i < 64 ? z1 : z2), i & 31);
          const float x = zi * dg[i];
          if (lane == (i & 31)) bx[i] = x;
          z0 = fmaf(-l0, x, z0); z1 = fmaf(-l1, x, z1); z2 = fmaf(-l2, x, z2);
        }
      }
      __syncthreads();
      // depth back-substitution dZ = Q (u - E^T dX) and patch retraction for this CTA's patches
      for (int r = r_lo + threadIdx.x; r < r_hi; r += blockDim.x) {
        const float* Ek = a.Ecols + (size_t)r * n6;
        float s = 0.f;
        for (int c = 0; c < n6; c++) s += Ek[c] * bx[c];
        const float dz = a.qk[r] * (a.uk[r] - s);
        if (a.p.kx[r] >= a.n_patches) continue;
        float* pdz = a.patches + (size_t)a.p.kx[r] * 3 * PP + 2 * PP;
        float dd = pdz[0] + dz;
        dd = (dd > 20.f) ? 1.0f : dd;  // patch_retr_kernel, ba_cuda.cu:200-203
        dd = fmaxf(dd, 1e-4f);
        for (int t = 0; t < PP; t++) pdz[t] = dd;
      }
      if (blockIdx.x == 0 && threadIdx.x < N) {  // pose_retr_kernel, ba_cuda.cu:160-188
        float* Pp = a.poses + 7 * (size_t)(a.t0 + threadIdx.x);
        Vec3 t{Pp[0], Pp[1], Pp[2]};
        Quat q{Pp[3], Pp[4], Pp[5], Pp[6]};
        se3_retract(bx + 6 * threadIdx.x, t, q);
        Pp[0] = t.x; Pp[1] = t.y; Pp[2] = t.z; Pp[3] = q.x; Pp[4] = q.y; Pp[5] = q.z; Pp[6] = q.w;
      }
      grid.sync();
    } else {  // structure-only (ba_cuda.cu:494-506): dZ = Q u
      if (a.kk_check) {   // reused plan: agree on its validity before applying anything (see above)
        grid.sync();
        if (*reinterpret_cast<volatile int*>(a.status) & WV_STATUS_STALE_PLAN) break;
      }
      for (int r = r_lo + threadIdx.x; r < r_hi; r += blockDim.x) {
        if (a.p.kx[r] >= a.n_patches) continue;
        float* pdz = a.patches + (size_t)a.p.kx[r] * 3 * PP + 2 * PP;
        float dd = pdz[0] + a.qk[r] * a.uk[r];
        dd = (dd > 20.f) ? 1.0f : dd;
        dd = fmaxf(dd, 1e-4f);
        for (int t = 0; t < PP; t++) pdz[t] = dd;
      }
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// reproject: fused per-edge projection of all PxP pixels, no depth clamp (ba_cuda.cu:368-418)
// ---------------------------------------------------------------------------------------------------
__global__ void ba_reproject_kernel(const float* __restrict__ poses, const float* __restrict__ patches,
                                    const float* __restrict__ intr, const int64_t* __restrict__ ii,
                                    const int64_t* __restrict__ jj, const int64_t* __restrict__ kk, int64_t E,
                                    int64_t n_poses, int64_t n_patches, int P, float* __restrict__ coords, int* status) {
  const int PP = P * P;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int64_t total = E * PP;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = t / PP; const int q = (int)(t % PP);
    const int64_t ix = ii[n], jx = jj[n], kx = kk[n];
    float* o = coords + n * 2 * PP + q;
    if (ix < 0 || ix >= n_poses || jx < 0 || jx >= n_poses || kx < 0 || kx >= n_patches) {
      set_status(status, WV_STATUS_BAD_INDEX);
      o[0] = 0.f; o[PP] = 0.f;
      continue;
    }
    const float* Pi = poses + 7 * ix; const float* Pj = poses + 7 * jx;
    Vec3 tij; Quat qij;
    relative_pose({Pi[0], Pi[1], Pi[2]}, {Pi[3], Pi[4], Pi[5], Pi[6]}, {Pj[0], Pj[1], Pj[2]}, {Pj[3], Pj[4], Pj[5], Pj[6]}, tij, qij);
    const float* pc = patches + kx * 3 * PP + q;
    const Vec3 Xi{(pc[0] - cx) / fx, (pc[PP] - cy) / fy, 1.f};
    const float W = pc[2 * PP];
    const Vec3 Rx = rotate(qij, Xi);
    const float X = Rx.x + W * tij.x, Y = Rx.y + W * tij.y, Z = Rx.z + W * tij.z;
    o[0] = fx * (X / Z) + cx;
    o[PP] = fy * (Y / Z) + cy;
  }
}

int ba_reproject(const float* poses, const float* patches, const float* intr, const int64_t* ii, const int64_t* jj,
                 const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, float* coords, int* status,
                 cudaStream_t st) {
  if (E == 0) return WV_OK;
  const int64_t total = E * P * P;
  int64_t g = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (g > cap) g = cap;
  ba_reproject_kernel<<<(int)g, 256, 0, st>>>(poses, patches, intr, ii, jj, kk, E, n_poses, n_patches, P, coords, status);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

// ---------------------------------------------------------------------------------------------------
// host side of cuda_ba.forward (windowed path) and cuda_ba.neighbors
// ---------------------------------------------------------------------------------------------------
static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
int plan_grid();  // ba_plan.cu

size_t ba_workspace_bytes(int64_t E, int64_t n_patches, int N) {
  const int G = plan_grid();
  const int64_t n6 = 6 * (int64_t)N, nS = n6 * (n6 + 1) / 2 + n6;
  const int64_t mp_cap = E < n_patches ? E : n_patches;
  return plan_bytes(E, n_patches, G) + al256(sizeof(float) * (size_t)(G * nS + 1)) + al256(sizeof(float) * (size_t)(nS + 1)) +
         al256(sizeof(float) * (size_t)(mp_cap * n6 + 1)) + 2 * al256(sizeof(float) * (size_t)(mp_cap + 1));
}

int ba_forward_window(float* poses, float* patches, const float* intr, const float* target, const float* weight,
                      const float* lmbda, const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                      int64_t n_poses, int64_t n_patches, int P, int t0, int t1, int iterations,
                      const float* patches_est, float mu, void* ws, size_t ws_bytes, int* status, cudaStream_t st,
                      int64_t plan_groups) {
  // plan_groups > 0: `ws` already holds the plan wv_neighbors built for these (kk, jj) with n_groups == plan_groups
  const int N = t1 - t0;
  if (N > kMaxFree) return WV_EUNSUPPORTED;
  if (E == 0 || iterations <= 0) return WV_OK;
  const int64_t groups = plan_groups > 0 ? plan_groups : n_patches;
  if (ws_bytes < ba_workspace_bytes(E, groups, N)) return WV_EWORKSPACE;
  const int G = plan_grid();
  const int n6 = 6 * N, nS = n6 * (n6 + 1) / 2 + n6;
  BAArgs a{};
  char* b = (char*)ws;
  size_t o = plan_carve(&a.p, ws, E, groups, G);
  auto take = [&](size_t bytes) { void* r = b + o; o += al256(bytes); return r; };
  a.partial = (float*)take(sizeof(float) * (size_t)(G * nS + 1));
  a.Sfull = (float*)take(sizeof(float) * (size_t)(nS + 1));
  a.Ecols = (float*)take(sizeof(float) * (size_t)(a.p.Mp_cap * n6 + 1));
  a.qk = (float*)take(sizeof(float) * (size_t)(a.p.Mp_cap + 1));
  a.uk = (float*)take(sizeof(float) * (size_t)(a.p.Mp_cap + 1));
  a.poses = poses; a.patches = patches; a.intr = intr; a.target = target; a.weight = weight; a.lmbda = lmbda;
  a.ii = ii; a.jj = jj; a.patches_est = patches_est; a.mu = mu; a.P = P; a.t0 = t0; a.N = N; a.iterations = iterations;
  a.status = status; a.n_poses = (int)n_poses; a.n_patches = (int)n_patches;
  a.kk_check = plan_groups > 0 ? kk : nullptr;
  if (plan_groups <= 0) {
    int rc = plan_build(a.p, kk, jj, ii, n_poses, nullptr, nullptr, status, st);
    if (rc != WV_OK) return rc;
  }
  const size_t smem = sizeof(float) * std::max((size_t)(kBaWarps * nS + kBaWarps * n6 + 64), (size_t)((n6 + 1) * (n6 + 1) + 2 * n6 + 64));
  if (smem > 48 * 1024)
    WV_CUDA_TRY(cudaFuncSetAttribute(ba_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  void* args[] = {&a};
  int Gb = G;   // CTAs of the BA kernel itself (<= the plan's grid: the partial-sum array is sized for G)
  if (const char* e = getenv("WVB200_BA_GRID")) { const int v = atoi(e); if (v > 0 && v < G) Gb = v; }
  WV_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)ba_window_kernel, dim3(Gb), dim3(kBaThreads), args, smem, st));
  ++g_launches;
  return WV_OK;
}

// latency probe for the windowed BA's roofline model (bench.py roofline_ba): `iters` grid-wide barriers of an otherwise empty
// cooperative grid of the BA kernel's shape; the caller times the launch.
__global__ void __launch_bounds__(kBaThreads) gridsync_probe_kernel(int iters, int* sink) {
  cg::grid_group grid = cg::this_grid();
  int acc = 0;
  for (int i = 0; i < iters; ++i) {
    grid.sync();
    acc += i;
  }
  if (sink && acc == -1) *sink = acc;
}

int gridsync_probe(int n_ctas, int iters, cudaStream_t st) {
  if (n_ctas <= 0) n_ctas = plan_grid();
  int* sink = nullptr;
  void* args[] = {&iters, &sink};
  WV_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)gridsync_probe_kernel, dim3(n_ctas), dim3(kBaThreads), args, 0, st));
  ++g_launches;
  return WV_OK;
}

int ba_neighbors(const int64_t* g, const int64_t* jj, int64_t E, int64_t n_groups, int64_t* ix, int64_t* jx, void* ws,
                 size_t ws_bytes, int* status, cudaStream_t st) {
  if (E == 0) return WV_OK;
  const int G = plan_grid();
  if (ws_bytes < plan_bytes(E, n_groups, G)) return WV_EWORKSPACE;
  Plan p;
  plan_carve(&p, ws, E, n_groups, G);
  return plan_build(p, g, jj, nullptr, 0, ix, jx, status, st);
}

}  // namespace wv
