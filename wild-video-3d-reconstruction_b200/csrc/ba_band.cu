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
// whole life: warp w < T owns the block columns C == w (mod T), T = bs + 2, and its lane d owns block (C + d, C); one more warp
// (w == T) owns the right-hand side, lane l the rows R == l (mod T).  A step is
//     (1) issue the global loads of the blocks that enter the window (one or two steps before their first use; the damping
//         S_dd <- S_dd (1 + 1e-4) + 1 of ba_cuda.cu:517-518 is applied a step later, when the load has landed),
//     (2) trailing update   A(R, C) -= Y(R, p) Y(C, p)^T   of the live blocks as packed fp32 FMAs (fma.rn.f32x2: the same two IEEE
//         FMAs per instruction, half the issue slots -- the update phase is issue-bound), Y(., p) read from shared memory where it
//         is kept twice: negated with every element doubled (the row operand, one 16-byte load = two ready pairs) and plain (the
//         column operand, one address per warp: a broadcast),
//     (3) the warp of column p + 1 -- whose blocks just got their last update -- factors its diagonal block (6x6 Cholesky in
//         registers, rsqrt; every lane does it, interleaved with the solve of its own sub-diagonal block
//         Y(R, p + 1) L^T = A(R, p + 1)) and publishes Y(., p + 1) (shared memory, double buffered); the warp that idles in the
//         next step copies the panel from shared memory to the factor record in global memory,
//     (r) the right-hand-side warp: z_p = L_pp^-1 y_p in the lane of row p, then y_R -= Y(R, p) z_p in the others,
//     (4) ONE __syncthreads.
// (3) overlaps the other warps' (2), so the factorisation of the diagonal block is off the critical path.
// ba_band_backsolve_kernel: one warp, L^T x = z backwards; see there.
#include <cstdlib>

#include "wv_common.cuh"

namespace wv {

static constexpr int kBandMaxBs = 25;   // T + 1 = 28 warps = 896 threads, 72 registers per thread
static constexpr int kYStride = 36;     // floats between the transposed Y tiles in shared memory (9 float4: odd, so the lanes' 16-byte loads spread over the banks)
static constexpr int kBackPrefetch = 6; // records in flight ahead of the back substitution

int band_max_bs() { return kBandMaxBs; }
int band_ld(int bs) { return (6 * (bs + 2) + 3) & ~3; }                       // > 6 bs + 5 + 3, multiple of 4 (vector REDs)
size_t band_floats(int n6, int bs) { return ((size_t)n6 * (size_t)(band_ld(bs) + 1) + 64 + 3) & ~(size_t)3; }   // multiple of 4: y behind it stays 16-byte aligned
// factor storage: per pose ONE record of (bs + 2) tiles of 36 floats: [0] L_pp, [1..bs] Y(p + d, p)^T, [bs + 1] = the reciprocals
// of diag(L_pp) in floats 0..5 and z_p in floats 8..13
// ... followed by the scratch of the two-sided form: the middle system of bs poses (band buffer, right-hand side, its own factor)
size_t band_factor_floats(int n6, int bs) {
  const int bsm = bs - 1 < 2 ? 2 : bs - 1;
  return (size_t)(n6 / 6) * (size_t)(bs + 2) * 36 + band_floats(6 * bs, bsm) + (size_t)(6 * bs + 4) + (size_t)bs * (size_t)(bsm + 2) * 36 + 64;
}

struct BandArgs {
  const float* S;   // band buffer
  const float* y;   // [6N]
  float* F;         // factor records
  float* dX;        // [6N]
  int N, bs, ld, damp;
  int* status;
  int prio;         // 1: the publishing warp updates first, alone (named barrier 1).  WVB200_BAND_PRIO; measured slower, off
  // Two-sided ("twisted") factorisation, grid = 2: CTA 0 eliminates poses 0 .. mf-1 forwards, CTA 1 poses N-1 .. N-mb downwards (the
  // same code on the block-reversed matrix); the g = N - mf - mb = bs poses in between form the MIDDLE system, which both CTAs
  // accumulate into Smid / ymid (band layout, leading dimension ldm) and a single-sided solve finishes.  Single-sided: mf = N,
  // mb = 0, Smid = nullptr, grid = 1.
  int mf, mb;
  float* Smid;
  float* ymid;
  int ldm;
};


// WV_BAND_TRACE (tools/band_trace.cu only): clock() stamps of the publishing warp's phases, one row of 8 per step
#ifdef WV_BAND_TRACE
__device__ unsigned g_band_trace[4096][8];
#define WV_TR(step, slot, cond)                                                              \
  do {                                                                                       \
    if ((cond) && (step) >= 0 && (step) < 4096) g_band_trace[step][slot] = (unsigned)clock(); \
  } while (0)
#else
#define WV_TR(step, slot, cond) do { } while (0)
#endif

// element (i, j) of a block held as float2 a[6][3]; i, j are compile-time under full unrolling
#define WV_E(a, i, j) (((j) & 1) ? (a)[i][(j) >> 1].y : (a)[i][(j) >> 1].x)

__device__ __forceinline__ float rsqrt_raw(float x) {   // one MUFU.RSQ: no denormal rescue (a pivot is >= 1 after the damping)
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__global__ void __launch_bounds__(32 * (kBandMaxBs + 3), 1) ba_band_factor_kernel(BandArgs g) {
  extern __shared__ __align__(16) float smem[];
  const int bs = g.bs, T = bs + 2, ld = g.ld;
  // this CTA's side: it eliminates `me` poses of a local chain of N poses -- its own plus, two-sided, the middle ones; local pose R
  // is global pose R (forward) or g.N - 1 - R (reverse)
  const bool rev = blockIdx.x == 1;
  const int me = rev ? g.mb : g.mf;
  const int gsz = g.N - g.mf - g.mb;
  const int N = me + gsz;
  float* const F = g.F + (rev ? (size_t)g.mf * (size_t)(bs + 2) * 36 : 0);
  float* sY = smem;                               // [2][T][kYStride]: Y(R, p) transposed ([k][i]), slot R % T, double buffered over p
  float* sL = sY + 2 * T * kYStride;              // [2][44]: L_qq (36) and the reciprocals of its diagonal (6)
  float* sz = sL + 88;                            // [8]: z_p
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const size_t FP = (size_t)(bs + 2) * 36;
  const int p0 = -1;                              // step -1 publishes column 0
  int m = (w + 1) % T;                            // tile warps: (w - p) mod T, kept incrementally; column C = p + m
  int ml = (lane + 1) % T;                        // rhs warp: the same for lane l
  const int d = lane;
  // every block column a tile warp owns is == w (mod T), so the shared-memory slots of a thread's two operands never change
  const int slotC = w, slotR = (w + d >= T) ? w + d - T : w + d;
  float2 a[6][3];                                 // tile warps: the block; rhs warp: row 0 holds the six right-hand-side values
#define WV_Y(i) WV_E(a, 0, i)
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) a[i][j] = make_float2(0.f, 0.f);
  // block (C + d, C) of the local chain into t.  Forward: as stored.  Reverse: local (R, C) is global (N-1-R, N-1-C), which lies
  // in the upper triangle: the stored block (N-1-C, N-1-R) transposed -- except on the diagonal, where the stored lower triangle
  // is what is wanted.  (A diagonal block's upper triangle gets whatever lies next to the diagonal in the band buffer: the
  // updates write it, nothing reads it.)
  auto fetch_tile = [&](int C, float2 (&t)[6][3]) {
    const int gr = rev ? g.N - 1 - C : C + d, gc = rev ? g.N - 1 - (C + d) : C;
    const float* src = g.S + (size_t)(6 * gr) * ld + 6 * gc;
    if (rev && d > 0) {
#pragma unroll
      for (int i = 0; i < 6; i++)
#pragma unroll
        for (int j = 0; j < 6; j++) WV_E(t, i, j) = src[(size_t)j * ld + i];
    } else {
#pragma unroll
      for (int i = 0; i < 6; i++) {
        const float2* row = reinterpret_cast<const float2*>(src + (size_t)i * ld);
        t[i][0] = row[0]; t[i][1] = row[1]; t[i][2] = row[2];
      }
    }
  };
  auto load_tile = [&](int C) { fetch_tile(C, a); };
  auto damp_diag = [&](float2 (&t)[6][3]) {
#pragma unroll
    for (int i = 0; i < 6; i++) WV_E(t, i, i) = WV_E(t, i, i) + (WV_E(t, i, i) * 1e-4f + 1.0f);
  };
  auto y_row = [&](int R) { return g.y + 6 * (rev ? g.N - 1 - R : R); };
  if (w < T) {                                    // tile warp w starts with column w
    if (d <= bs && w + d < N) load_tile(w);
    if (d == 0 && g.damp) damp_diag(a);
  } else if (lane < T && lane < N) {              // rhs warp: rows 0..T-1
#pragma unroll
    for (int i = 0; i < 6; i++) WV_Y(i) = y_row(lane)[i];
  }

  for (int p = p0; p < me; p++) {
    if (w < T) {
      // ---- (1) the warp whose column was panel p - 1 is free: it loads its next column, p + T, whole -- a warp-uniform branch (a
      //          load in SOME lane would put its destination registers on the scoreboard of the whole warp and stall the update),
      //          two steps before the column's first use, which is the damping of its diagonal block --------------------------------
      if (m == 0 && p >= 0) {
        // ... after it has written its factor record to global memory for the back substitution -- from shared memory, where panel p
        // stays valid through this step: the TRANSPOSED blocks Y(p + d, p)^T, and L_pp with the reciprocals of its diagonal
        float* Fp = F + (size_t)p * FP;
        if (d >= 1 && d <= bs && p + d < N) {
          const float4* src = reinterpret_cast<const float4*>(sY + (size_t)((p & 1) * T + slotR) * kYStride);
          float4* dst = reinterpret_cast<float4*>(Fp + (size_t)d * 36);
#pragma unroll
          for (int t = 0; t < 9; t++) dst[t] = src[t];
        } else if (d == 0 || d == 31) {
          const float4* src = reinterpret_cast<const float4*>(sL + (p & 1) * 44);
          float4* dst = reinterpret_cast<float4*>(Fp + (d == 0 ? 0 : (size_t)(bs + 1) * 36));
          if (d == 0) {
#pragma unroll
            for (int t = 0; t < 9; t++) dst[t] = src[t];
          } else {
            const float4 r0 = src[9], r1 = src[10];
            dst[0] = r0; dst[1] = make_float4(r1.x, r1.y, 0.f, 0.f);
          }
        }
        if (d <= bs && p + T + d < N) load_tile(p + T);
      }
      if (d == 0 && m == bs && g.damp && p + m >= T) damp_diag(a);
      // ---- (2) trailing update with panel p.  The warp that publishes in this step is the critical path (update -> factor + solve
      //          -> publish -> barrier).  g.prio lets it update first, alone (named barrier 1 over the tile warps), with the others'
      //          update hidden behind its factor + solve; measured (tools/band_trace.cu) that loses: 2.66 vs 2.21 us per step -----------
      WV_TR(p, 0, m == 1 && lane == 0);
      WV_TR(p, 6, m == 5 && lane == 0);
      if (g.prio && m != 1) asm volatile("bar.sync 1, %0;" ::"r"(32 * T) : "memory");
      if (p >= 0 && m >= 1 && d <= bs - m && p + m + d < N) {
        const float4* YR = reinterpret_cast<const float4*>(sY + (size_t)((p & 1) * T + slotR) * kYStride);   // this lane's row
        const float4* YC = reinterpret_cast<const float4*>(sY + (size_t)((p & 1) * T + slotC) * kYStride);   // the warp's column: a broadcast
#pragma unroll
        for (int h = 0; h < 3; h++) {   // two k per round: 12 floats = 3 float4 of each operand
          const float4 r0 = YR[3 * h], r1 = YR[3 * h + 1], r2 = YR[3 * h + 2];
          const float4 c0 = YC[3 * h], c1 = YC[3 * h + 1], c2 = YC[3 * h + 2];
          const float ra[6] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y}, rb[6] = {r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};
          const float2 ca[3] = {make_float2(c0.x, c0.y), make_float2(c0.z, c0.w), make_float2(c1.x, c1.y)};
          const float2 cb[3] = {make_float2(c1.z, c1.w), make_float2(c2.x, c2.y), make_float2(c2.z, c2.w)};
#pragma unroll
          for (int i = 0; i < 6; i++)   // make_float2(-x, -x): FFMA2 takes a negated scalar operand (-R.F32), no extra instruction
#pragma unroll
            for (int j = 0; j < 3; j++) a[i][j] = __ffma2_rn(make_float2(-ra[i], -ra[i]), ca[j], a[i][j]);
#pragma unroll
          for (int i = 0; i < 6; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) a[i][j] = __ffma2_rn(make_float2(-rb[i], -rb[i]), cb[j], a[i][j]);
        }
      }
      WV_TR(p, 1, m == 1 && lane == 0);
      WV_TR(p, 7, m == 5 && lane == 0);
      if (g.prio && m == 1) asm volatile("bar.arrive 1, %0;" ::"r"(32 * T) : "memory");
      // ---- (3) the warp of column q = p + 1: factor, solve, publish ----------------------------------------------------------------
      if (m == 1 && p + 1 >= 0 && p + 1 < me) {
        const int q = p + 1;
        float* L = sL + (q & 1) * 44;
        // lane 0 hands its diagonal block to the whole warp, and EVERY lane factors it: in lockstep that costs what lane 0 alone
        // would, and each lane's own forward substitution  Y L^T = A  (column j needs row j of L, final after factor step j) fills
        // the latency bubbles of the factorisation's serial chain -- instead of waiting for it to end.
        if (d == 0) {
#pragma unroll
          for (int i = 0; i < 6; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) *reinterpret_cast<float2*>(L + 6 * i + 2 * j) = a[i][j];
        }
        __syncwarp();
        float2 gd[6][3];
        {
          const float4* L4 = reinterpret_cast<const float4*>(L);
#pragma unroll
          for (int t = 0; t < 9; t++) {
            const float4 v = L4[t];
            gd[(2 * t) / 3][(2 * t) % 3] = make_float2(v.x, v.y);
            gd[(2 * t + 1) / 3][(2 * t + 1) % 3] = make_float2(v.z, v.w);
          }
        }
        __syncwarp();   // lane 0 overwrites the buffer with L below
        float dmin = WV_E(gd, 0, 0);
        float rd[6];
#pragma unroll
        for (int j = 0; j < 6; j++) {
          const float dj = WV_E(gd, j, j);
          dmin = fminf(dmin, dj);
          const float r = rsqrt_raw(dj);
          rd[j] = r;
          WV_E(gd, j, j) = dj * r;
#pragma unroll
          for (int i = j + 1; i < 6; i++) WV_E(gd, i, j) *= r;
#pragma unroll
          for (int k = j + 1; k < 6; k++)
#pragma unroll
            for (int i = k; i < 6; i++) WV_E(gd, i, k) = fmaf(-WV_E(gd, i, j), WV_E(gd, k, j), WV_E(gd, i, k));
          // column j of this lane's block:  Y[i][j] = (A[i][j] - sum_{k < j} Y[i][k] L[j][k]) / L[j][j]   (lane 0's result is not used)
#pragma unroll
          for (int k = 0; k < j; k++)
#pragma unroll
            for (int i = 0; i < 6; i++) WV_E(a, i, j) = fmaf(-WV_E(a, i, k), WV_E(gd, j, k), WV_E(a, i, j));
#pragma unroll
          for (int i = 0; i < 6; i++) WV_E(a, i, j) *= r;
        }
        WV_TR(p, 2, lane == 0);
        if (d == 0) {   // L_qq and the reciprocals of its diagonal: for the right-hand-side warp and the factor record (next step)
          if (!(dmin > 0.f)) set_status(g.status, WV_STATUS_NOT_SPD);   // a non-positive pivot (or NaN)
#define WV_LZ(f) ((f) % 6 <= (f) / 6 ? WV_E(gd, (f) / 6, (f) % 6) : 0.f)   /* flat element f of L, zero above the diagonal */
          float4* L4 = reinterpret_cast<float4*>(L);
#pragma unroll
          for (int t = 0; t < 9; t++) L4[t] = make_float4(WV_LZ(4 * t), WV_LZ(4 * t + 1), WV_LZ(4 * t + 2), WV_LZ(4 * t + 3));
          L4[9] = make_float4(rd[0], rd[1], rd[2], rd[3]);
          L4[10] = make_float4(rd[4], rd[5], 0.f, 0.f);
#undef WV_LZ
        }
        WV_TR(p, 3, lane == 1);
        if (d >= 1 && d <= bs && q + d < N) {
          float4* Yt = reinterpret_cast<float4*>(sY + (size_t)((q & 1) * T + slotR) * kYStride);   // transposed: [k][i] = Y[i][k]
#pragma unroll
          for (int t = 0; t < 9; t++)
            Yt[t] = make_float4(WV_E(a, (4 * t) % 6, (4 * t) / 6), WV_E(a, (4 * t + 1) % 6, (4 * t + 1) / 6),
                                WV_E(a, (4 * t + 2) % 6, (4 * t + 2) / 6), WV_E(a, (4 * t + 3) % 6, (4 * t + 3) / 6));
        }
      }
      WV_TR(p, 4, m == 1 && lane == 1);
      m = (m == 0) ? T - 1 : m - 1;
    } else if (p >= 0) {
      // ---- (r) right-hand side: lane l holds y of the rows == l (mod T) ------------------------------------------------------------
      if (lane < T && ml == 0) {   // row p: every update is in; L z_p = y_p
        const float* L = sL + (p & 1) * 44;
        float* zg = F + (size_t)p * FP + (size_t)(bs + 1) * 36 + 8;
        float zz[6];
#pragma unroll
        for (int i = 0; i < 6; i++) {
          float s = WV_Y(i);
#pragma unroll
          for (int k = 0; k < i; k++) s = fmaf(-L[6 * i + k], zz[k], s);
          zz[i] = s * L[36 + i];
          sz[i] = zz[i];
          zg[i] = zz[i];
        }
        const int Rn = p + T;      // this lane's next row
#pragma unroll
        for (int i = 0; i < 6; i++) WV_Y(i) = Rn < N ? y_row(Rn)[i] : 0.f;
      }
      __syncwarp();
      if (lane < T && ml >= 1 && ml <= bs && p + ml < N) {
        const float4* YC = reinterpret_cast<const float4*>(sY + (size_t)((p & 1) * T + lane) * kYStride);
        float yt[36];
#pragma unroll
        for (int t = 0; t < 9; t++) { const float4 v = YC[t]; yt[4 * t] = v.x; yt[4 * t + 1] = v.y; yt[4 * t + 2] = v.z; yt[4 * t + 3] = v.w; }
#pragma unroll
        for (int k = 0; k < 6; k++) {
          const float zk = sz[k];
#pragma unroll
          for (int i = 0; i < 6; i++) WV_Y(i) = fmaf(-yt[6 * k + i], zk, WV_Y(i));
        }
      }
    }
    if (w == T) ml = (ml == 0) ? T - 1 : ml - 1;
    WV_TR(p, 5, w == T && lane == 0);
    __syncthreads();
  }
  // ---- two-sided: what is left in the window is this side's Schur complement on the middle poses (local me .. me + gsz - 1): block
  //      (me + m + d, me + m) in the thread of phase m (the phases now belong to p = me).  Middle = W_forward + W_reverse - A (the
  //      original block is in both windows, damped in both): the forward side subtracts it.  The reverse side's blocks land
  //      block-reversed and transposed.
  if (g.Smid) {
    const int ldm = g.ldm;
    if (w < T) {
      const int cM = m, rM = m + d;
      if (m < gsz && d <= bs && rM < gsz) {
        if (!rev) {
          float2 o[6][3];
          fetch_tile(me + m, o);
          if (d == 0 && g.damp) damp_diag(o);
#pragma unroll
          for (int i = 0; i < 6; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) { a[i][j].x -= o[i][j].x; a[i][j].y -= o[i][j].y; }
        }
        const int rb = rev ? gsz - 1 - cM : rM, cb = rev ? gsz - 1 - rM : cM;
        float* dst = g.Smid + (size_t)(6 * rb) * ldm + 6 * cb;
#pragma unroll
        for (int i = 0; i < 6; i++)
#pragma unroll
          for (int j = 0; j < 6; j++) {
            if (d == 0) { if (j <= i) atomicAdd(dst + (size_t)i * ldm + j, WV_E(a, i, j)); }
            else if (!rev) atomicAdd(dst + (size_t)i * ldm + j, WV_E(a, i, j));
            else atomicAdd(dst + (size_t)j * ldm + i, WV_E(a, i, j));
          }
      }
    } else if (lane < T && ml < gsz) {
      const int R = me + ml, rb = rev ? gsz - 1 - ml : ml;
#pragma unroll
      for (int i = 0; i < 6; i++) atomicAdd(g.ymid + 6 * rb + i, WV_Y(i) - (rev ? 0.f : y_row(R)[i]));
    }
  }
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// L^T x = z, backwards, by one warp, "push" style: when x_R is known (lane 0: six-step substitution with L_RR), every lane d adds
// Y(R, R - d)^T x_R to the running sum of pose R - d that it carries, and the sums move down one lane -- so that lane 0 holds the
// complete sum of pose R - 1 when the next step begins.  The chain of a step is substitution -> 6 shuffles -> 36 FMAs -> 6 shuffles;
// no reduction tree.  The factor records (one contiguous run of (bs + 2) * 36 floats per pose) stream through a cp.async ring in
// shared memory that holds the bs + 1 records a step reads plus kBackPrefetch in flight.
__global__ void __launch_bounds__(32, 1) ba_band_backsolve_kernel(BandArgs g) {
  extern __shared__ __align__(16) float smem[];
  const int bs = g.bs, d = threadIdx.x;
  // this CTA's side (see ba_band_factor_kernel): `me` eliminated poses with factor records, then gsz middle poses whose x is
  // already in dX (the middle solve ran before this kernel)
  const bool rev = blockIdx.x == 1;
  const int me = rev ? g.mb : g.mf;
  const int gsz = g.N - g.mf - g.mb;
  const int N = me + gsz;
  const int FP = (bs + 2) * 36, FP4 = FP / 4, RING = bs + 2 + kBackPrefetch;
  const float* const F = g.F + (rev ? (size_t)g.mf * FP : 0);
  float* ring = smem;                                   // [RING][FP], record of pose q in slot q % RING
  // record `pose` into ring slot `slot` (= pose % RING, kept incrementally by the callers: no division in the loop)
  auto issue = [&](int pose, int slot) {
    if (pose >= 0 && pose < me) {
      const float4* src = reinterpret_cast<const float4*>(F) + (size_t)pose * FP4 + d;
      float4* dst = reinterpret_cast<float4*>(ring) + (size_t)slot * FP4 + d;
#pragma unroll
      for (int t = 0; t < (kBandMaxBs + 2) * 9; t += 32)
        if (t + d < FP4) cp_async16(dst + t, src + t);
    }
    cp_async_commit();                                  // an empty group keeps the count uniform
  };
  {
    int sl = (N - 1) % RING;
    for (int s = 0; s <= bs + kBackPrefetch; s++) { issue(N - 1 - s, sl); sl = sl == 0 ? RING - 1 : sl - 1; }
  }
  float acc[6];
#pragma unroll
  for (int j = 0; j < 6; j++) acc[j] = 0.f;
  const bool tile_lane = d >= 1 && d <= bs;
  int slot0 = (N - 1) % RING;                           // R % RING
  int slotd = ((N - 1 - d) % RING + RING) % RING;       // (R - d) % RING
  for (int R = N - 1; R >= 0; R--) {
    cp_async_wait<kBackPrefetch>();                     // record R - bs (and every later one) has landed
    __syncwarp();
    issue(R - bs - kBackPrefetch - 1, slot0 + 1 == RING ? 0 : slot0 + 1);   // the slot of pose R + 1, which nobody reads any more
    const bool act = tile_lane && R - d >= 0 && R - d < me;
    float yt[36];                                       // Y(R, R - d)^T: [j][i]
    if (act) {
      const float4* t4 = reinterpret_cast<const float4*>(ring + (size_t)slotd * FP + d * 36);
#pragma unroll
      for (int q = 0; q < 9; q++) { const float4 v = t4[q]; yt[4 * q] = v.x; yt[4 * q + 1] = v.y; yt[4 * q + 2] = v.z; yt[4 * q + 3] = v.w; }
    }
    float x[6];
#pragma unroll
    for (int k = 0; k < 6; k++) x[k] = 0.f;
    if (d == 0) {
      float* xg = g.dX + 6 * (rev ? g.N - 1 - R : R);
      if (R >= me) {   // a middle pose: solved already
#pragma unroll
        for (int k = 0; k < 6; k++) x[k] = xg[k];
      } else {         // L_RR^T x_R = z_R - sum:  x[k] = (rhs[k] - sum_{j > k} L[j][k] x[j]) / L[k][k], k descending
        const float* L = ring + (size_t)slot0 * FP;
        const float* rz = L + (bs + 1) * 36;            // [0..5] reciprocals of the diagonal, [8..13] z_R
#pragma unroll
        for (int k = 5; k >= 0; k--) {
          float s = rz[8 + k] - acc[k];
#pragma unroll
          for (int j = k + 1; j < 6; j++) s = fmaf(-L[6 * j + k], x[j], s);
          x[k] = s * rz[k];
          xg[k] = x[k];
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 6; k++) x[k] = __shfl_sync(0xffffffffu, x[k], 0);
    if (act) {
#pragma unroll
      for (int j = 0; j < 6; j++)
#pragma unroll
        for (int i = 0; i < 6; i++) acc[j] = fmaf(yt[6 * j + i], x[i], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < 6; j++) acc[j] = __shfl_down_sync(0xffffffffu, acc[j], 1);   // lanes > bs carry zeros
    slot0 = slot0 == 0 ? RING - 1 : slot0 - 1;
    slotd = slotd == 0 ? RING - 1 : slotd - 1;
  }
}

static int band_mid_bs(int bs) { return bs - 1 < 2 ? 2 : bs - 1; }   // half-width of the middle system of bs poses
static bool band_two_sided(int N, int bs) {
  static const int on = [] { const char* e = getenv("WVB200_BAND_TWISTED"); return e ? atoi(e) : 1; }();
  return on && N >= 2 * bs + 16;
}

static int band_launch(const BandArgs& a, int grid, cudaStream_t st) {
  const int T = a.bs + 2;
  const size_t smem = sizeof(float) * (size_t)(2 * T * kYStride + 88 + 8);
  ba_band_factor_kernel<<<grid, 32 * (T + 1), smem, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}
static int band_back_launch(const BandArgs& a, int grid, cudaStream_t st) {
  const int T = a.bs + 2;
  const size_t smem_b = sizeof(float) * (size_t)(a.bs + 2 + kBackPrefetch) * T * 36;   // 128 KB at bs = 25
  WV_CUDA_TRY(cudaFuncSetAttribute(ba_band_backsolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b));
  ba_band_backsolve_kernel<<<grid, 32, smem_b, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

// S: band buffer (band_floats(n6, bs) floats, lower triangle, not yet damped); y [n6]; F: band_factor_floats(n6, bs) floats of
// scratch; dX [n6].  Launches on `st` only, no host synchronisation; a non-positive pivot sets WV_STATUS_NOT_SPD.
// Long chains are factored from BOTH ends at once (two CTAs, N / 2 dependent steps each instead of N): see BandArgs.
int ba_band_solve(const float* S, const float* y, int n6, int bs, int damp, float* F, float* dX, int* status, cudaStream_t st) {
  if (n6 <= 0) return WV_OK;
  if (bs < 2 || bs > kBandMaxBs || n6 % 6 != 0) return WV_EINVAL;
  const int N = n6 / 6;
  static const int prio = [] { const char* e = getenv("WVB200_BAND_PRIO"); return e ? atoi(e) : 0; }();
  BandArgs a{S, y, F, dX, N, bs, band_ld(bs), damp, status, prio, N, 0, nullptr, nullptr, 0};
  if (!band_two_sided(N, bs)) {
    int rc = band_launch(a, 1, st);
    return rc != WV_OK ? rc : band_back_launch(a, 1, st);
  }
  const int g = bs, bsm = band_mid_bs(bs);
  a.mf = (N - g + 1) / 2;
  a.mb = N - g - a.mf;
  a.ldm = band_ld(bsm);
  float* mid = F + (size_t)N * (size_t)(bs + 2) * 36;
  const size_t smid = band_floats(6 * g, bsm), ymid = (size_t)(6 * g + 3) & ~(size_t)3;
  a.Smid = mid;
  a.ymid = mid + smid;
  float* Fmid = a.ymid + ymid;
  WV_CUDA_TRY(cudaMemsetAsync(mid, 0, (smid + ymid) * sizeof(float), st));
  int rc = band_launch(a, 2, st);
  if (rc != WV_OK) return rc;
  // the middle system: bs poses, already damped, solved single-sided straight into its place in dX
  BandArgs am{a.Smid, a.ymid, Fmid, dX + 6 * a.mf, g, bsm, a.ldm, 0, status, prio, g, 0, nullptr, nullptr, 0};
  if ((rc = band_launch(am, 1, st)) != WV_OK) return rc;
  if ((rc = band_back_launch(am, 1, st)) != WV_OK) return rc;
  return band_back_launch(a, 2, st);
}

}  // namespace wv
