// Band Cholesky solve of the global BA's reduced camera system when it is BANDED -- the usual case: a patch is observed by
// frames within +-PATCH_LIFETIME of its source frame (dpvo/dpvo.py:756-769) or by the fixed pattern of
// global_bundle_adjustment (dpvo/dpvo.py:448-458: i+1 and i+10..i+19), so S couples two free poses only when they are at most
// `span` frames apart.  The reference factorises the dense 6N x 6N matrix whatever its structure (torch::linalg::cholesky +
// cholesky_solve, dpvo/fastba/ba_cuda.cu:521-523: 6.2e11 flop at N = 2047); the band needs N * span^2 * 6^3 / 2 (1.4e8 at
// span = 24) -- but as a chain of N dependent steps, so the design goal is a short step, not throughput.
//
// Band layout (band_ld / band_floats): element (r, c) of the lower triangle, 0 <= r - c < ld, lives at S[r * ld + c] -- a
// dense-looking addressing whose rows slide by one column per row -- so ba_global_kernel<true> builds straight into it
// (GArgs::ld) and an all-reduce over ranks moves 6N (ld + 1) floats instead of (6N)^2 (15 MB instead of 603 MB at N = 2047).
//
// ba_band_factor_kernel: ONE CTA walks the chain, one step per pose p ("panel" = its 6 columns).  The live part of the matrix is
// the lower triangle of a (bs + 1) x (bs + 1) window of 6x6 pose blocks; every block lives in the REGISTERS of one thread for its
// whole life: warp w owns the block columns C == w (mod T), T = bs + 2, lane d owns block (C + d, C), lane 31 the right-hand
// side of row C.  A step is
//     (1) issue the global loads of the block row that enters the window (row p + bs + 1; damping applied to its diagonal block),
//     (2) trailing update   A(R, C) -= Y(R, p) Y(C, p)^T   for the live blocks, Y(., p) read from shared memory,
//     (3) the warp of column p + 1 -- whose blocks just got their last update -- factors its diagonal block in lane 0
//         (6x6 Cholesky in registers), hands L to its other lanes through shared memory, which solve their sub-diagonal
//         blocks Y(R, p + 1) L^T = A(R, p + 1), forward-substitutes its right-hand side, and publishes Y(., p + 1) (shared
//         memory, double buffered) and the factor (global memory),
//     (4) ONE __syncthreads.
// (3) overlaps the other warps' (2), so the factorisation of the diagonal block is off the critical path.
// ba_band_backsolve_kernel: one warp, L^T x = z backwards, lane d reduces block (p + d, p); the next step's blocks are
// prefetched.  The damping S_dd <- S_dd (1 + 1e-4) + 1 (ba_cuda.cu:517-518) is applied when a diagonal block is loaded.
#include "wv_common.cuh"

namespace wv {

static constexpr int kBandMaxBs = 26;   // T = 28 warps = 896 threads, 73 registers per thread
static constexpr int kTileStride = 38;  // floats between transposed 6x6 tiles in shared memory (8-byte aligned, bank spread)

int band_max_bs() { return kBandMaxBs; }
int band_ld(int bs) { return (6 * (bs + 2) + 3) & ~3; }                       // > 6 bs + 5 + 3, multiple of 4 (vector REDs)
size_t band_floats(int n6, int bs) { return (size_t)n6 * (size_t)(band_ld(bs) + 1) + 64; }
// factor storage: per pose (bs + 2) tiles of 36 floats: [0] L_pp, [1..bs] Y(p + d, p), [bs + 1] the reciprocals of diag(L_pp); then z [n6]
size_t band_factor_floats(int n6, int bs) { return (size_t)(n6 / 6) * (size_t)(bs + 2) * 36 + (size_t)n6 + 64; }

struct BandArgs {
  const float* S;   // band buffer
  const float* y;   // [6N]
  float* F;         // factor tiles
  float* z;         // [6N] forward-substituted right-hand side
  float* dX;        // [6N]
  int N, bs, ld, damp;
  int* status;
};

// in-place 6x6 Cholesky of the lower triangle of a (upper part ignored, zeroed on exit); rd = reciprocals of the diagonal of L.
// rsqrtf (2 ulp) instead of sqrtf + division: this runs in ONE lane on the critical path of every step.  false on a bad pivot.
__device__ __forceinline__ bool chol6(float (&a)[6][6], float (&rd)[6]) {
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 6; j++) {
    float d = a[j][j];
    if (!(d > 0.f)) { ok = false; d = 1.f; }
    const float r = rsqrtf(d);
    a[j][j] = d * r;
    rd[j] = r;
#pragma unroll
    for (int i = j + 1; i < 6; i++) a[i][j] *= r;
#pragma unroll
    for (int k = j + 1; k < 6; k++)
#pragma unroll
      for (int i = k; i < 6; i++) a[i][k] = fmaf(-a[i][j], a[k][j], a[i][k]);
  }
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j < 6; j++)
      if (j > i) a[i][j] = 0.f;
  return ok;
}

__global__ void __launch_bounds__(32 * (kBandMaxBs + 2), 1) ba_band_factor_kernel(BandArgs g) {
  extern __shared__ float smem[];
  const int bs = g.bs, T = bs + 2, N = g.N, ld = g.ld;
  float* sYt = smem;                              // [2][T][kTileStride]: Y(R, p) transposed ([k][i]), slot R % T
  float* sz = smem + 2 * T * kTileStride;         // [2][8]: z_p
  float* sL = sz + 16;                            // [36 + 6]: L_qq and the reciprocals of its diagonal, for the lanes of the publishing warp
  const int lane = threadIdx.x & 31, cm = threadIdx.x >> 5;
  const int d = lane;
  const bool is_tile = d <= bs, is_rhs = lane == 31;
  const size_t FP = (size_t)(bs + 2) * 36;
  // every block column this warp owns is == cm (mod T), so the shared-memory slots of this thread's rows never change
  const int slotC = cm, slotR = (cm + d >= T) ? cm + d - T : cm + d;
  float a[6][6];
  float yv[6];
#pragma unroll
  for (int i = 0; i < 6; i++) {
    yv[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 6; j++) a[i][j] = 0.f;
  }
  const int p0 = -(bs + 1);
  int m = ((cm - p0) % T + T) % T;                // (cm - p) mod T, kept incrementally

  for (int p = p0; p < N; p++) {
    // ---- (1) loads of the entering blocks: the diagonal block of row p + bs + 1, the other blocks of row p + bs + 2 (one step
    //          before their first use, so that the load latency never meets the publishing warp) ---------------------------------
    {
      const int R = p + bs + 1 + (d > 0 ? 1 : 0);
      if (is_tile && m == bs + 1 - d + (d > 0 ? 1 : 0) && R < N) {
        const int C = R - d;
        if (C >= 0) {
#pragma unroll
          for (int i = 0; i < 6; i++) {
            const float2* row = reinterpret_cast<const float2*>(g.S + (size_t)(6 * R + i) * ld + 6 * C);
            const float2 v0 = row[0], v1 = row[1], v2 = row[2];
            a[i][0] = v0.x; a[i][1] = v0.y; a[i][2] = v1.x; a[i][3] = v1.y; a[i][4] = v2.x; a[i][5] = v2.y;
          }
          if (d == 0) {
#pragma unroll
            for (int i = 0; i < 6; i++) {
              if (g.damp) a[i][i] = a[i][i] + (a[i][i] * 1e-4f + 1.0f);
#pragma unroll
              for (int j = 0; j < 6; j++) if (j > i) a[i][j] = 0.f;
            }
          }
        }
      }
      if (is_rhs && m == bs + 1) {
        const int Ry = p + bs + 1;
#pragma unroll
        for (int i = 0; i < 6; i++) yv[i] = Ry < N ? g.y[6 * Ry + i] : 0.f;
      }
    }
    // ---- (2) trailing update with panel p --------------------------------------------------------------------------------------
    if (p >= 0) {
      const float* Yb = sYt + (size_t)(p & 1) * T * kTileStride;
      if (is_tile && m >= 1 && m <= bs - d) {
        const int C = p + m, R = C + d;
        if (R < N) {
          const float* YR = Yb + slotR * kTileStride;
          const float* YC = Yb + slotC * kTileStride;
#pragma unroll
          for (int k = 0; k < 6; k++) {
            const float2 r0 = *reinterpret_cast<const float2*>(YR + 6 * k), r1 = *reinterpret_cast<const float2*>(YR + 6 * k + 2),
                         r2 = *reinterpret_cast<const float2*>(YR + 6 * k + 4);
            const float2 c0 = *reinterpret_cast<const float2*>(YC + 6 * k), c1 = *reinterpret_cast<const float2*>(YC + 6 * k + 2),
                         c2 = *reinterpret_cast<const float2*>(YC + 6 * k + 4);
            const float yr[6] = {r0.x, r0.y, r1.x, r1.y, r2.x, r2.y};
            const float yc[6] = {c0.x, c0.y, c1.x, c1.y, c2.x, c2.y};
#pragma unroll
            for (int i = 0; i < 6; i++)
#pragma unroll
              for (int j = 0; j < 6; j++) a[i][j] = fmaf(-yr[i], yc[j], a[i][j]);
          }
        }
      }
      if (is_rhs && m >= 1 && m <= bs) {
        const int R = p + m;
        if (R < N) {
          const float* YR = Yb + slotC * kTileStride;
          const float* zp = sz + (p & 1) * 8;
#pragma unroll
          for (int k = 0; k < 6; k++) {
            const float zk = zp[k];
#pragma unroll
            for (int i = 0; i < 6; i++) yv[i] = fmaf(-YR[6 * k + i], zk, yv[i]);
          }
        }
      }
    }
    // ---- (3) the warp of column q = p + 1: factor, scale, publish -----------------------------------------------------------------
    if (m == 1 && p + 1 >= 0 && p + 1 < N) {
      const int q = p + 1;
      if (d == 0) {
        float rd[6];
        if (!chol6(a, rd)) set_status(g.status, WV_STATUS_NOT_SPD);
#pragma unroll
        for (int i = 0; i < 6; i++) {
          sL[36 + i] = rd[i];
#pragma unroll
          for (int j = 0; j < 6; j++) sL[6 * i + j] = a[i][j];
        }
      }
      __syncwarp();
      const int R = q + d;
      float* Fq = g.F + (size_t)q * FP;
      if (is_tile && R < N) {
        if (d > 0) {   // Y L^T = A by forward substitution along each row:  Y[i][j] = (A[i][j] - sum_{k < j} Y[i][k] L[j][k]) / L[j][j]
#pragma unroll
          for (int j = 0; j < 6; j++) {
#pragma unroll
            for (int k = 0; k < j; k++) {
              const float c = sL[6 * j + k];
#pragma unroll
              for (int i = 0; i < 6; i++) a[i][j] = fmaf(-a[i][k], c, a[i][j]);
            }
            const float r = sL[36 + j];
#pragma unroll
            for (int i = 0; i < 6; i++) a[i][j] *= r;
          }
          float* Yt = sYt + (size_t)(q & 1) * T * kTileStride + slotR * kTileStride;
#pragma unroll
          for (int k = 0; k < 6; k++)
#pragma unroll
            for (int i = 0; i < 6; i++) Yt[6 * k + i] = a[i][k];
        }
        float4* dst = reinterpret_cast<float4*>(Fq + (size_t)d * 36);
#pragma unroll
        for (int t = 0; t < 9; t++) {
          const int e = 4 * t;
          dst[t] = make_float4(a[e / 6][e % 6], a[(e + 1) / 6][(e + 1) % 6], a[(e + 2) / 6][(e + 2) % 6], a[(e + 3) / 6][(e + 3) % 6]);
        }
        if (d == 0) {   // reciprocals of the diagonal, for the back substitution
          float4* di = reinterpret_cast<float4*>(Fq + (size_t)(bs + 1) * 36);
          di[0] = make_float4(sL[36], sL[37], sL[38], sL[39]);
          di[1] = make_float4(sL[40], sL[41], 0.f, 0.f);
        }
      }
      if (is_rhs) {   // L z_q = y_q
        float* zq = sz + (q & 1) * 8;
        float zz[6];
#pragma unroll
        for (int i = 0; i < 6; i++) {
          float s = yv[i];
#pragma unroll
          for (int k = 0; k < i; k++) s = fmaf(-sL[6 * i + k], zz[k], s);
          zz[i] = s * sL[36 + i];
          zq[i] = zz[i];
          g.z[6 * q + i] = zz[i];
        }
      }
    }
    __syncthreads();
    m = (m == 0) ? T - 1 : m - 1;
  }
}

// L^T x = z, backwards, by one warp: lane d = 1..bs holds block Y(p + d, p), lane 0 holds L_pp and (next) the reciprocals of its diagonal; the blocks of step p - 1
// are loaded while step p is reduced.
__global__ void __launch_bounds__(32, 1) ba_band_backsolve_kernel(BandArgs g) {
  __shared__ float sx[kBandMaxBs + 2][8];
  const int bs = g.bs, T = bs + 2, N = g.N, d = threadIdx.x;
  const size_t FP = (size_t)(bs + 2) * 36;
  float cur[36], nxt[36], zc[6], zn[6], rc[6], rn[6];
  auto fetch = [&](int p, float* t, float* zz, float* rr) {
    if (d <= bs && (d == 0 || p + d < N)) {
      const float4* src = reinterpret_cast<const float4*>(g.F + (size_t)p * FP + (size_t)d * 36);
#pragma unroll
      for (int q = 0; q < 9; q++) { const float4 v = src[q]; t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w; }
    }
    if (d == 0) {
      const float4* rs = reinterpret_cast<const float4*>(g.F + (size_t)p * FP + (size_t)(bs + 1) * 36);
      const float4 r0 = rs[0], r1 = rs[1];
      rr[0] = r0.x; rr[1] = r0.y; rr[2] = r0.z; rr[3] = r0.w; rr[4] = r1.x; rr[5] = r1.y;
#pragma unroll
      for (int i = 0; i < 6; i++) zz[i] = g.z[6 * p + i];
    }
  };
#pragma unroll
  for (int i = 0; i < 36; i++) { cur[i] = 0.f; nxt[i] = 0.f; }
#pragma unroll
  for (int i = 0; i < 6; i++) { zc[i] = 0.f; zn[i] = 0.f; rc[i] = 0.f; rn[i] = 0.f; }
  if (N > 0) fetch(N - 1, cur, zc, rc);
  for (int p = N - 1; p >= 0; p--) {
    if (p > 0) fetch(p - 1, nxt, zn, rn);
    float part[6];
#pragma unroll
    for (int j = 0; j < 6; j++) part[j] = 0.f;
    if (d >= 1 && d <= bs && p + d < N) {
      const float* xR = sx[(p + d) % T];
#pragma unroll
      for (int i = 0; i < 6; i++) {
        const float xi = xR[i];
#pragma unroll
        for (int j = 0; j < 6; j++) part[j] = fmaf(cur[6 * i + j], xi, part[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 6; j++) part[j] = warp_sum(part[j]);
    if (d == 0) {   // L_pp^T x_p = z_p - sum:  x[k] = (rhs[k] - sum_{j > k} L[j][k] x[j]) / L[k][k], k descending
      float x[6];
#pragma unroll
      for (int k = 5; k >= 0; k--) {
        float s = zc[k] - part[k];
#pragma unroll
        for (int j = k + 1; j < 6; j++) s = fmaf(-cur[6 * j + k], x[j], s);
        x[k] = s * rc[k];
        sx[p % T][k] = x[k];
        g.dX[6 * p + k] = x[k];
      }
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 36; i++) cur[i] = nxt[i];
#pragma unroll
    for (int i = 0; i < 6; i++) { zc[i] = zn[i]; rc[i] = rn[i]; }
  }
}

// S: band buffer (band_floats(n6, bs) floats, lower triangle, not yet damped); y [n6]; F: band_factor_floats(n6, bs) floats of
// scratch; dX [n6].  Two launches on `st`, no host synchronisation; a non-positive pivot sets WV_STATUS_NOT_SPD.
int ba_band_solve(const float* S, const float* y, int n6, int bs, int damp, float* F, float* dX, int* status, cudaStream_t st) {
  if (n6 <= 0) return WV_OK;
  if (bs < 2 || bs > kBandMaxBs || n6 % 6 != 0) return WV_EINVAL;
  const int N = n6 / 6, T = bs + 2;
  BandArgs a{S, y, F, F + (size_t)N * (size_t)(bs + 2) * 36, dX, N, bs, band_ld(bs), damp, status};
  const size_t smem = sizeof(float) * (size_t)(2 * T * kTileStride + 16 + 44);
  ba_band_factor_kernel<<<1, 32 * T, smem, st>>>(a);
  WV_LAUNCH_CHECK();
  ba_band_backsolve_kernel<<<1, 32, 0, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

}  // namespace wv
