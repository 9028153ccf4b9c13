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
static constexpr int kMaxEntries = 384;  // (frame, 6-vector) entries of one patch's column of E (per-warp list in shared memory)
static constexpr size_t kEntryBytes = (size_t)kGWarps * kMaxEntries * (6 * sizeof(float) + sizeof(int));
// A patch with more than kMaxEntries / 2 edges ("heavy": a camera turning on the spot under distance-based edges can see one
// patch from every frame; the reference has no limit) may not fit the shared-memory list.  The main pass skips it and appends its
// rank to a list; a second, small launch of the SAME kernel body (HEAVY) walks that list with the entry lists in global memory
// (kHeavyCap entries per warp: two per pose of a 4 096-frame buffer).  Slower per patch, but nothing is dropped.
static constexpr int kHeavyCtas = 16;
static constexpr int kHeavyCap = 2 * 4096 + 64;
static constexpr size_t kHeavySpillBytes = (size_t)kHeavyCtas * kGWarps * kHeavyCap * (6 * sizeof(float) + sizeof(int));
static size_t heavy_list_ints(int64_t E) { return (size_t)(E / (kMaxEntries / 2 + 1) + 2 + 3) & ~(size_t)3; }

struct GArgs {
  float* poses; float* patches; const float* intr; const float* target; const float* weight; const float* lmbda;
  const int64_t* ii; const int64_t* jj;
  const float* patches_est; float mu;
  int P, t0, N, n_poses;
  Plan p;
  float* S;         // element (r, c), r >= c, at S[r * ld + c]: dense [6N][6N] (ld = 6N) or the band layout of ba_band.cu
  int ld;
  float* y;         // [6N]
  const float* dX;  // [N][6]
  float* dZ;        // [Mp_cap], indexed by the plan's patch rank
  int shard, nshards;
  int* status;
  int* heavy;       // [0] count, [4 ...] ranks of the heavy patches of this pass (workspace, behind the plan)
  float* spill;     // HEAVY: entry lists in global memory, [kHeavyCtas * kGWarps][kHeavyCap] x (6 floats), then the frames
};

// A CTA works on a chunk of kChunk consecutive patches (consecutive ids = the same or neighbouring source frames, which observe
// nearly the same set of frames), 8 consecutive patches per warp.
//   * The Schur update of the chunk, sum_k q_k e_k e_k^T restricted to the diagonal window of frames [wb, wb + kWinFrames) that
//     (almost) all of its entries fall into, is a small SYRK: the window parts of the columns e_k are laid out densely in shared
//     memory (Ecol [kChunk][192]) and every thread accumulates 8x8 register tiles of the window's lower triangle over the 64
//     columns -- no atomics -- and adds its tiles to S in HBM once per chunk.  Pairs with an entry outside the window
//     (loop-closure edges) go to HBM as atomics directly.
//   * The pose-block terms of an edge ((j,j), (i,j) and v_j: 63 values) are run-length accumulated in the lane's registers while
//     consecutive patches of the warp keep the same (i, j) in that lane, and flushed with RED atomics when the key changes;
//     the source-frame terms ((i,i), v_i) likewise per warp.
static constexpr int kChunk = 64;
static constexpr int kPerWarp = kChunk / kGWarps;  // 8 consecutive patches per warp
static constexpr int kWinFrames = 32;
static constexpr int kWinDim = 6 * kWinFrames;  // 192
static constexpr int kTile = 8;
static constexpr int kTilesPerDim = kWinDim / kTile;                    // 24
static constexpr int kTiles = kTilesPerDim * (kTilesPerDim + 1) / 2;    // 300 lower tiles
static constexpr size_t kBuildSmem = kEntryBytes + sizeof(float) * (kChunk * kWinDim + 2 * kChunk);

// 8-byte vector RED (sm_90+): rows of a 6x6 block start at an even float offset (n6 and 6*fb are even), S is 256-B aligned
__device__ __forceinline__ void red2(float* p, float x, float y) { atomicAdd(reinterpret_cast<float2*>(p), make_float2(x, y)); }

// block (fa, fb) of S with fa >= fb += s A B^T, straight to HBM (fa == fb: lower triangle of the block only)
__device__ __forceinline__ void red_block(float* S, int ld, int fa, int fb, const float* A, const float* B, float s) {
  float* base = S + (size_t)(6 * fa) * ld + 6 * fb;
  if (fa != fb) {
#pragma unroll
    for (int r = 0; r < 6; r++) {
      const float ar = s * A[r];
      red2(base + (size_t)r * ld, ar * B[0], ar * B[1]);
      red2(base + (size_t)r * ld + 2, ar * B[2], ar * B[3]);
      red2(base + (size_t)r * ld + 4, ar * B[4], ar * B[5]);
    }
  } else {
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
      for (int c = 0; c <= r; c++) atomicAdd(base + (size_t)r * ld + c, s * A[r] * B[c]);
  }
}

struct LaneAcc {  // run-length accumulators of one lane: pose-block terms of edges (fi -> fj)
  int fi, fj;     // free-pose indices of the current key (may be out of [0, N): fixed pose)
  float JJ[21], IJ[36], vj[6];
};

__device__ __forceinline__ void lane_acc_reset(LaneAcc& A, int fi, int fj) {
  A.fi = fi; A.fj = fj;
#pragma unroll
  for (int t = 0; t < 21; t++) A.JJ[t] = 0.f;
#pragma unroll
  for (int t = 0; t < 36; t++) A.IJ[t] = 0.f;
#pragma unroll
  for (int t = 0; t < 6; t++) A.vj[t] = 0.f;
}

__device__ __forceinline__ void lane_acc_flush(const LaneAcc& A, float* S, float* y, int N, int ld) {
  const bool free_i = A.fi >= 0 && A.fi < N, free_j = A.fj >= 0 && A.fj < N;
  if (!free_j) return;
  {
    float* b = S + (size_t)(6 * A.fj) * ld + 6 * A.fj;
    int t = 0;
#pragma unroll
    for (int p = 0; p < 6; p++) {
#pragma unroll
      for (int c = 0; c <= p; c++) atomicAdd(b + (size_t)p * ld + c, A.JJ[t++]);
      atomicAdd(y + 6 * A.fj + p, A.vj[p]);
    }
  }
  if (!free_i) return;
  if (A.fi > A.fj) {
    float* b = S + (size_t)(6 * A.fi) * ld + 6 * A.fj;
#pragma unroll
    for (int p = 0; p < 6; p++)
#pragma unroll
      for (int c = 0; c < 6; c += 2) red2(b + (size_t)p * ld + c, -A.IJ[6 * p + c], -A.IJ[6 * p + c + 1]);
  } else if (A.fi < A.fj) {
    float* b = S + (size_t)(6 * A.fj) * ld + 6 * A.fi;
#pragma unroll
    for (int c = 0; c < 6; c++)
#pragma unroll
      for (int p = 0; p < 6; p += 2) red2(b + (size_t)c * ld + p, -A.IJ[6 * p + c], -A.IJ[6 * (p + 1) + c]);
  } else {
    float* b = S + (size_t)(6 * A.fi) * ld + 6 * A.fi;
#pragma unroll
    for (int p = 0; p < 6; p++)
#pragma unroll
      for (int c = 0; c <= p; c++) atomicAdd(b + (size_t)p * ld + c, -(A.IJ[6 * p + c] + A.IJ[6 * c + p]));
  }
}

template <bool BUILD, bool HEAVY>
__global__ void __launch_bounds__(kGThreads) ba_global_kernel(GArgs a) {
  extern __shared__ float smem_all[];  // entry lists [kGWarps][kMaxEntries] x (6 floats + frame); BUILD: + Ecol, q, qu
  float(*ent_v)[kMaxEntries][6] = reinterpret_cast<float(*)[kMaxEntries][6]>(smem_all);
  int(*ent_f)[kMaxEntries] = reinterpret_cast<int(*)[kMaxEntries]>(smem_all + kGWarps * kMaxEntries * 6);
  float* dyn_mem = smem_all + kEntryBytes / sizeof(float);  // BUILD: Ecol [kChunk][kWinDim], q [kChunk], qu [kChunk]
  constexpr int kCap = HEAVY ? kHeavyCap : kMaxEntries;
  __shared__ int s_wb;
  __shared__ int perm[kGWarps][32];
  const int N = a.N, n6 = 6 * N;
  const int ld = a.ld;  // row stride of S: n6 (dense) or the band layout's (ba_band.cu)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int PP = a.P * a.P, ctr = (a.P / 2) * a.P + a.P / 2;
  const float fx = a.intr[0], fy = a.intr[1], cx = a.intr[2], cy = a.intr[3];  // row 0 only (ba_cuda.cu:232-238)
  const float lambda = a.lmbda[0];
  const int Mp = a.p.hdr->Mp;
  // the main pass walks this shard's contiguous run of patch ranks; the heavy pass walks the list the main pass left behind
  const int r_lo = HEAVY ? 0 : (int)((long long)Mp * a.shard / a.nshards);
  const int r_hi = HEAVY ? a.heavy[0] : (int)((long long)Mp * (a.shard + 1) / a.nshards);
  const unsigned FULL = 0xffffffffu;
  const size_t gw = (size_t)blockIdx.x * kGWarps + warp;
  float(*ev)[6] = HEAVY ? reinterpret_cast<float(*)[6]>(a.spill + gw * kHeavyCap * 6) : ent_v[warp];
  int* ef = HEAVY ? reinterpret_cast<int*>(a.spill + (size_t)kHeavyCtas * kGWarps * kHeavyCap * 6) + gw * kHeavyCap : ent_f[warp];
  float* Ecol = dyn_mem;
  float* qs = dyn_mem + kChunk * kWinDim;
  float* qus = qs + kChunk;
  const int nchunks = (r_hi - r_lo + kChunk - 1) / kChunk;

  for (int ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
    const int c_lo = r_lo + ch * kChunk, c_hi = min(c_lo + kChunk, r_hi);
    int wb = 0;
    if (BUILD) {
      // window base: the lowest free frame the chunk's first patch touches (its segment is ordered by target frame)
      if (threadIdx.x == 0) {
        const int e0 = a.p.eperm[a.p.seg_off[HEAVY ? a.heavy[4 + c_lo] : c_lo]];
        long long f = min(a.jj[e0], a.ii[e0]) - a.t0;
        s_wb = (int)max(0ll, min(f, (long long)max(N - kWinFrames, 0)));
      }
      for (int t = threadIdx.x; t < kChunk * kWinDim + 2 * kChunk; t += kGThreads) dyn_mem[t] = 0.f;
      __syncthreads();
      wb = s_wb;
    }
    LaneAcc acc;
    float II[21], vi[6];
    int acc_f0 = -1;  // free index of the source frame the II / vi accumulators belong to (warp-uniform), -1: none
    if (BUILD) {
      lane_acc_reset(acc, -1, -1);
#pragma unroll
      for (int t = 0; t < 21; t++) II[t] = 0.f;
#pragma unroll
      for (int t = 0; t < 6; t++) vi[t] = 0.f;
    }
    for (int pi = 0; pi < kPerWarp; pi++) {
      const int slot = c_lo + warp * kPerWarp + pi;     // position in the chunk's run (column slot - c_lo of Ecol)
      if (slot >= c_hi) break;
      const int r = HEAVY ? a.heavy[4 + slot] : slot;   // rank of the patch in the plan
      const int id = a.p.kx[r], off = a.p.seg_off[r], n = a.p.seg_off[r + 1] - off;
      if (!HEAVY && 2 * n > kMaxEntries) {              // may not fit the shared-memory list: leave it to the heavy pass
        if (lane == 0) a.heavy[4 + atomicAdd(a.heavy, 1)] = r;
        continue;
      }
      const float* pc = a.patches + (size_t)id * 3 * PP;
      const float px = pc[ctr], py = pc[PP + ctr], pd = pc[2 * PP + ctr];
      float Csum = 0.f, usum = 0.f;
      int L = 0;  // entries so far (warp-uniform)
      bool overflow = false;
      for (int base = 0; base < n; base += 32) {
        bool act = base + lane < n;
        int e = act ? a.p.eperm[off + base + lane] : 0;
        int ix = -1, jx = -1;
        if (act) {
          const int64_t iv = a.ii[e], jv = a.jj[e];
          if (iv < 0 || iv >= a.n_poses || jv < 0 || jv >= a.n_poses) { set_status(a.status, WV_STATUS_BAD_INDEX); act = false; }
          else { ix = (int)iv; jx = (int)jv; }
        }
        if (BUILD) {
          // lane assignment: edges into the window first, in target-frame order, then the others (loop closures).  Patches of one
          // source frame then present the same (i, j) to the same lane, whatever loop edges they carry, and the run-length
          // accumulators below survive from patch to patch.
          const int fj0 = jx - a.t0;
          const bool inw = act && fj0 >= wb && fj0 < wb + kWinFrames;
          const unsigned m_in = __ballot_sync(FULL, inw), m_out = __ballot_sync(FULL, act && !inw);
          const unsigned below = (1u << lane) - 1u;
          const int np = inw ? __popc(m_in & below)
                             : (act ? __popc(m_in) + __popc(m_out & below) : __popc(m_in) + __popc(m_out) + __popc(~(m_in | m_out) & below));
          perm[warp][np] = lane;
          __syncwarp();
          const int src = perm[warp][lane];
          __syncwarp();
          e = __shfl_sync(FULL, e, src); ix = __shfl_sync(FULL, ix, src); jx = __shfl_sync(FULL, jx, src);
          act = __shfl_sync(FULL, (int)act, src) != 0;
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
        // source frame shared by all lanes of the chunk (the tracker's edge lists): its terms are reduced in the warp
        const int ix0 = __shfl_sync(FULL, ix, 0);
        const bool uni = __all_sync(FULL, !act || ix == ix0);
        const int f0 = ix0 - a.t0;
        const bool free0 = uni && f0 >= 0 && f0 < N;
        if (BUILD) {
          // ---- (j,j), (i,j), v_j: run-length accumulation in this lane
          if (act && (free_j || free_i)) {
            if (acc.fi != fi || acc.fj != fj) { lane_acc_flush(acc, a.S, a.y, N, ld); lane_acc_reset(acc, fi, fj); }
            int t = 0;
#pragma unroll
            for (int p = 0; p < 6; p++) {
#pragma unroll
              for (int c = 0; c <= p; c++) acc.JJ[t++] += T.w[0] * T.Jj[0][p] * T.Jj[0][c] + T.w[1] * T.Jj[1][p] * T.Jj[1][c];
#pragma unroll
              for (int c = 0; c < 6; c++) acc.IJ[6 * p + c] += T.w[0] * T.Ji[0][p] * T.Jj[0][c] + T.w[1] * T.Ji[1][p] * T.Jj[1][c];
              acc.vj[p] += T.w[0] * T.r[0] * T.Jj[0][p] + T.w[1] * T.r[1] * T.Jj[1][p];
            }
          }
          // ---- (i,i), v_i
          const int want = free0 ? f0 : -1;
          if (acc_f0 != want && acc_f0 >= 0) {  // warp-uniform: reduce + flush the previous source frame
            int t = 0;
#pragma unroll
            for (int p = 0; p < 6; p++) {
#pragma unroll
              for (int c = 0; c <= p; c++) {
                const float v = warp_sum(II[t]);
                if (lane == t) atomicAdd(a.S + (size_t)(6 * acc_f0 + p) * ld + 6 * acc_f0 + c, v);
                II[t] = 0.f;
                t++;
              }
              const float wv = warp_sum(vi[p]);
              if (lane == 21 + p) atomicAdd(a.y + 6 * acc_f0 + p, -wv);
              vi[p] = 0.f;
            }
          }
          acc_f0 = want;
          if (free0) {
            if (act) {
              int t = 0;
#pragma unroll
              for (int p = 0; p < 6; p++) {
#pragma unroll
                for (int c = 0; c <= p; c++) II[t++] += T.w[0] * T.Ji[0][p] * T.Ji[0][c] + T.w[1] * T.Ji[1][p] * T.Ji[1][c];
                vi[p] += T.w[0] * T.r[0] * T.Ji[0][p] + T.w[1] * T.r[1] * T.Ji[1][p];
              }
            }
          } else if (free_i) {  // mixed source frames inside one chunk of lanes: per-edge atomics
            red_block(a.S, ld, fi, fi, T.Ji[0], T.Ji[0], T.w[0]);
            red_block(a.S, ld, fi, fi, T.Ji[1], T.Ji[1], T.w[1]);
#pragma unroll
            for (int p = 0; p < 6; p++) atomicAdd(a.y + 6 * fi + p, -(T.w[0] * T.r[0] * T.Ji[0][p] + T.w[1] * T.r[1] * T.Ji[1][p]));
          }
        }
        // ---- entries of the patch's column of E
        {
          const unsigned mj = __ballot_sync(FULL, free_j);
          const int nj = __popc(mj);
          if (L + nj <= kCap) {
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
              if (L + 1 <= kCap) {
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
            if (L + ni <= kCap) {
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
          // window part of the column -> Ecol (entries of one patch may repeat a frame: add), the rest -> HBM atomics
          const int col = slot - c_lo;
          float* mycol = Ecol + col * kWinDim;
          for (int t = lane; t < L * 6; t += 32) {
            const int en = t / 6, c = t - 6 * en, f = ef[en];
            if (f >= wb && f < wb + kWinFrames) atomicAdd(mycol + 6 * (f - wb) + c, ev[en][c]);
            else atomicAdd(a.y + 6 * f + c, -q * u * ev[en][c]);
          }
          if (lane == 0) { qs[col] = q; qus[col] = q * u; }
          // pairs with an entry outside the window (loop-closure edges): one block per lane, straight to HBM
          bool any_out = false;
          for (int t = lane; t < L; t += 32) any_out |= ef[t] < wb || ef[t] >= wb + kWinFrames;
          if (__any_sync(FULL, any_out)) {
            for (int ea = 0; ea < L; ea++) {
              const int fa = ef[ea];
              if (fa >= wb && fa < wb + kWinFrames) continue;  // warp-uniform
              for (int eb = lane; eb < L; eb += 32) {
                const int fb = ef[eb];
                const bool in_b = fb >= wb && fb < wb + kWinFrames;
                if (fa > fb) red_block(a.S, ld, fa, fb, ev[ea], ev[eb], -q);
                else if (fa < fb) { if (in_b) red_block(a.S, ld, fb, fa, ev[eb], ev[ea], -q); }  // (out, out) is visited from the other side
                else red_block(a.S, ld, fa, fa, ev[ea], ev[eb], -q);
              }
            }
          }
        }
      } else {
        float sdot = 0.f;
        for (int t = lane; t < L * 6; t += 32) sdot += ev[t / 6][t % 6] * a.dX[6 * ef[t / 6] + t % 6];
        sdot = warp_sum(sdot);
        if (lane == 0) a.dZ[r] = q * (u - sdot);
      }
      __syncwarp();
    }
    if (BUILD) {
      // flush the run-length accumulators
      lane_acc_flush(acc, a.S, a.y, N, ld);
      if (acc_f0 >= 0) {
        int t = 0;
#pragma unroll
        for (int p = 0; p < 6; p++) {
#pragma unroll
          for (int c = 0; c <= p; c++) {
            const float v = warp_sum(II[t]);
            if (lane == t) atomicAdd(a.S + (size_t)(6 * acc_f0 + p) * ld + 6 * acc_f0 + c, v);
            t++;
          }
          const float wv = warp_sum(vi[p]);
          if (lane == 21 + p) atomicAdd(a.y + 6 * acc_f0 + p, -wv);
        }
      }
      __syncthreads();
      // ---- window SYRK: W -= sum_k q_k e_k e_k^T over the chunk's columns, 8x8 register tiles of the lower triangle
      const int wb6 = 6 * wb;
      const int ncol = c_hi - c_lo;
      for (int tile = threadIdx.x; tile < kTiles; tile += kGThreads) {
        int tr = (int)((sqrtf(8.f * tile + 1.f) - 1.f) * 0.5f);
        while ((tr + 1) * (tr + 2) / 2 <= tile) tr++;
        while (tr * (tr + 1) / 2 > tile) tr--;
        const int tc = tile - tr * (tr + 1) / 2;
        if (wb6 + tr * kTile >= n6) continue;
        float accT[kTile][kTile];
#pragma unroll
        for (int i = 0; i < kTile; i++)
#pragma unroll
          for (int j = 0; j < kTile; j++) accT[i][j] = 0.f;
        for (int k = 0; k < ncol; k++) {
          const float4* ra = reinterpret_cast<const float4*>(Ecol + k * kWinDim + tr * kTile);
          const float4* rb = reinterpret_cast<const float4*>(Ecol + k * kWinDim + tc * kTile);
          const float4 a0 = ra[0], a1 = ra[1], b0 = rb[0], b1 = rb[1];
          const float nq = -qs[k];
          const float av[8] = {a0.x * nq, a0.y * nq, a0.z * nq, a0.w * nq, a1.x * nq, a1.y * nq, a1.z * nq, a1.w * nq};
          const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
          for (int i = 0; i < kTile; i++)
#pragma unroll
            for (int j = 0; j < kTile; j++) accT[i][j] = fmaf(av[i], bv[j], accT[i][j]);
        }
#pragma unroll
        for (int i = 0; i < kTile; i++) {
          const int R = wb6 + tr * kTile + i;
          if (R >= n6) break;
          float* grow = a.S + (size_t)R * ld + wb6 + tc * kTile;
          if (tr != tc && ((wb6 + tc * kTile) & 3) == 0 && (ld & 3) == 0) {  // 16-B aligned rows: two vector REDs
            if (accT[i][0] != 0.f || accT[i][1] != 0.f || accT[i][2] != 0.f || accT[i][3] != 0.f)
              atomicAdd(reinterpret_cast<float4*>(grow), make_float4(accT[i][0], accT[i][1], accT[i][2], accT[i][3]));
            if (accT[i][4] != 0.f || accT[i][5] != 0.f || accT[i][6] != 0.f || accT[i][7] != 0.f)
              atomicAdd(reinterpret_cast<float4*>(grow + 4), make_float4(accT[i][4], accT[i][5], accT[i][6], accT[i][7]));
          } else {
#pragma unroll
            for (int j = 0; j < kTile; j++) {
              // lower triangle of S only (like every other writer)
              if (wb6 + tc * kTile + j <= R && accT[i][j] != 0.f) atomicAdd(grow + j, accT[i][j]);
            }
          }
        }
      }
      for (int t = threadIdx.x; t < kWinDim; t += kGThreads) {
        if (wb6 + t >= n6) break;
        float s = 0.f;
        for (int k = 0; k < ncol; k++) s = fmaf(-qus[k], Ecol[k * kWinDim + t], s);
        if (s != 0.f) atomicAdd(a.y + wb6 + t, s);
      }
      __syncthreads();
    }
  }
}

// dZ_k = q_k (u_k - sum_a E_a . dX[f_a])  (ba_cuda.cu:524-525) without the entry lists: the dot products are taken edge by edge
// (E_j . dX[j] + E_i . dX[i]) as the edge's terms are recomputed.  One warp per patch of this shard, a lane per edge; no shared
// memory, so several CTAs per SM (the build kernel's body with BUILD = false carried the 86 KB entry lists and 249 registers
// along for a sum it does not need: 0.72 ms at configs[3]; patches of any observation count take the same path).
__global__ void __launch_bounds__(kGThreads) ba_global_dz_kernel(GArgs a) {
  const int N = a.N, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int PP = a.P * a.P, ctr = (a.P / 2) * a.P + a.P / 2;
  const float fx = a.intr[0], fy = a.intr[1], cx = a.intr[2], cy = a.intr[3];  // row 0 only (ba_cuda.cu:232-238)
  const float lambda = a.lmbda[0];
  const int Mp = a.p.hdr->Mp;
  const int r_lo = (int)((long long)Mp * a.shard / a.nshards), r_hi = (int)((long long)Mp * (a.shard + 1) / a.nshards);
  for (int r = r_lo + blockIdx.x * kGWarps + warp; r < r_hi; r += gridDim.x * kGWarps) {
    const int id = a.p.kx[r], off = a.p.seg_off[r], n = a.p.seg_off[r + 1] - off;
    const float* pc = a.patches + (size_t)id * 3 * PP;
    const float px = pc[ctr], py = pc[PP + ctr], pd = pc[2 * PP + ctr];
    float Csum = 0.f, usum = 0.f, sdot = 0.f;
    for (int base = 0; base < n; base += 32) {
      if (base + lane >= n) continue;
      const int e = a.p.eperm[off + base + lane];
      const int64_t iv = a.ii[e], jv = a.jj[e];
      if (iv < 0 || iv >= a.n_poses || jv < 0 || jv >= a.n_poses) { set_status(a.status, WV_STATUS_BAD_INDEX); continue; }
      const int ix = (int)iv, jx = (int)jv, fi = ix - a.t0, fj = jx - a.t0;
      EdgeTerms T;
      edge_terms(a, e, ix, jx, px, py, pd, fx, fy, cx, cy, T);
      Csum += T.w[0] * T.Jz[0] * T.Jz[0] + T.w[1] * T.Jz[1] * T.Jz[1];
      usum += T.w[0] * T.r[0] * T.Jz[0] + T.w[1] * T.r[1] * T.Jz[1];
      if (fj >= 0 && fj < N) {
#pragma unroll
        for (int c = 0; c < 6; c++) sdot += (T.w[0] * T.Jz[0] * T.Jj[0][c] + T.w[1] * T.Jz[1] * T.Jj[1][c]) * a.dX[6 * fj + c];
      }
      if (fi >= 0 && fi < N) {
#pragma unroll
        for (int c = 0; c < 6; c++) sdot -= (T.w[0] * T.Jz[0] * T.Ji[0][c] + T.w[1] * T.Jz[1] * T.Ji[1][c]) * a.dX[6 * fi + c];
      }
    }
    float C = warp_sum(Csum), u = warp_sum(usum);
    sdot = warp_sum(sdot);
    if (a.patches_est) {  // depth prior (dpvo/ba.py:151-159)
      const float dest = a.patches_est[(size_t)id * 3 * PP + 2 * PP + ctr];
      if (dest > 0.f) { C += a.mu; u -= a.mu * (pd - dest); }
    }
    if (lane == 0) a.dZ[r] = (u - sdot) / (C + lambda);
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
// ba_band.cu: band layout of the reduced system for a half-width of `bs` frames
int band_ld(int bs);
size_t band_floats(int n6, int bs);

static size_t heavy_offset(int64_t E, int64_t n_patches) { return (plan_bytes(E, n_patches, plan_grid()) + 1024 + 255) & ~(size_t)255; }
size_t ba_global_workspace_bytes(int64_t E, int64_t n_patches) {
  return heavy_offset(E, n_patches) + heavy_list_ints(E) * sizeof(int) + 256 + kHeavySpillBytes;
}

static GArgs make_gargs(float* poses, float* patches, const float* intr, const float* target, const float* weight, const float* lmbda,
                        const int64_t* ii, const int64_t* jj, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0, int t1,
                        const float* patches_est, float mu, void* ws, int shard, int nshards, int* status) {
  GArgs a{};
  plan_carve(&a.p, ws, E, n_patches, plan_grid());
  a.poses = poses; a.patches = patches; a.intr = intr; a.target = target; a.weight = weight; a.lmbda = lmbda; a.ii = ii; a.jj = jj;
  a.patches_est = patches_est; a.mu = mu; a.P = P; a.t0 = t0; a.N = t1 - t0; a.n_poses = (int)n_poses;
  a.shard = shard; a.nshards = nshards; a.status = status;
  a.heavy = reinterpret_cast<int*>(static_cast<char*>(ws) + heavy_offset(E, n_patches));
  a.spill = reinterpret_cast<float*>(static_cast<char*>(ws) + ((heavy_offset(E, n_patches) + heavy_list_ints(E) * sizeof(int) + 255) & ~(size_t)255));
  return a;
}

static int global_grid(int64_t n_patches_guess) {
  const int64_t want = (n_patches_guess + kChunk - 1) / kChunk;
  const int64_t cap = (int64_t)sm_count() * 2;
  return (int)std::max<int64_t>(1, std::min(want, cap));
}

// the main pass over this shard's patches, then the heavy pass over the patches it skipped (an empty list costs one idle launch)
template <bool BUILD>
static int launch_global_passes(const GArgs& a, int64_t n_patches_guess, size_t smem, cudaStream_t st) {
  WV_CUDA_TRY(cudaMemsetAsync(a.heavy, 0, 4 * sizeof(int), st));
  WV_CUDA_TRY(cudaFuncSetAttribute(ba_global_kernel<BUILD, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ba_global_kernel<BUILD, false><<<global_grid(n_patches_guess), kGThreads, smem, st>>>(a);
  WV_LAUNCH_CHECK();
  WV_CUDA_TRY(cudaFuncSetAttribute(ba_global_kernel<BUILD, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ba_global_kernel<BUILD, true><<<kHeavyCtas, kGThreads, smem, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int ba_global_build(float* poses, float* patches, const float* intr, const float* target, const float* weight, const float* lmbda,
                    const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P,
                    int t0, int t1, const float* patches_est, float mu, int shard, int nshards, int build_plan, float* S, float* y,
                    void* ws, size_t ws_bytes, int* status, cudaStream_t st, int band_bs) {
  if (ws_bytes < ba_global_workspace_bytes(E, n_patches)) return WV_EWORKSPACE;
  GArgs a = make_gargs(poses, patches, intr, target, weight, lmbda, ii, jj, E, n_poses, n_patches, P, t0, t1, patches_est, mu, ws,
                       shard, nshards, status);
  const size_t n6 = 6 * (size_t)a.N;
  // band_bs > 0: S is the band buffer of ba_band.cu: element (r, c), 0 <= r - c < ld, lives at S[r * ld + c]
  const size_t s_floats = band_bs > 0 ? band_floats((int)n6, band_bs) : n6 * n6;
  a.ld = band_bs > 0 ? band_ld(band_bs) : (int)n6;
  a.S = S; a.y = y;
  if (build_plan) {
    const int rc = plan_build(a.p, kk, jj, ii, n_poses, nullptr, nullptr, status, st);
    if (rc != WV_OK) return rc;
  }
  if (n6) {
    WV_CUDA_TRY(cudaMemsetAsync(S, 0, s_floats * sizeof(float), st));
    WV_CUDA_TRY(cudaMemsetAsync(y, 0, n6 * sizeof(float), st));
  }
  return launch_global_passes<true>(a, std::min(E, n_patches), kBuildSmem, st);
}

// widest frame span of any patch's observations among the free poses (source frame included): the reduced system couples two
// free poses only if some patch sees both, so S is block-banded with a half-width of `span` pose blocks.  One warp per patch.
__global__ void __launch_bounds__(kGThreads) ba_global_span_kernel(GArgs a, int* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int Mp = a.p.hdr->Mp, N = a.N;
  int worst = 0;
  for (int r = blockIdx.x * kGWarps + warp; r < Mp; r += gridDim.x * kGWarps) {
    const int off = a.p.seg_off[r], n = a.p.seg_off[r + 1] - off;
    int lo = 1 << 30, hi = -1;
    for (int t = lane; t < n; t += 32) {
      const int e = a.p.eperm[off + t];
      const long long fi = a.ii[e] - a.t0, fj = a.jj[e] - a.t0;
      if (fi >= 0 && fi < N) { lo = min(lo, (int)fi); hi = max(hi, (int)fi); }
      if (fj >= 0 && fj < N) { lo = min(lo, (int)fj); hi = max(hi, (int)fj); }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
    if (hi >= lo) worst = max(worst, hi - lo);
  }
  if (lane == 0 && worst > 0) atomicMax(out, worst);
}

int ba_global_span(const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int t0, int t1,
                   int build_plan, int* span_out, void* ws, size_t ws_bytes, int* status, cudaStream_t st) {
  if (ws_bytes < ba_global_workspace_bytes(E, n_patches)) return WV_EWORKSPACE;
  GArgs a = make_gargs(nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, ii, jj, E, n_poses, n_patches, 3, t0, t1, nullptr, 0.f, ws, 0, 1, status);
  if (build_plan) {
    const int rc = plan_build(a.p, kk, jj, ii, n_poses, nullptr, nullptr, status, st);
    if (rc != WV_OK) return rc;
  }
  WV_CUDA_TRY(cudaMemsetAsync(span_out, 0, sizeof(int), st));
  ba_global_span_kernel<<<global_grid(std::min(E, n_patches) / 4 + 1), kGThreads, 0, st>>>(a, span_out);
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
  {
    const int64_t want = (std::min(E, n_patches) + kGWarps - 1) / kGWarps;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(want, (int64_t)sm_count() * 8));
    ba_global_dz_kernel<<<grid, kGThreads, 0, st>>>(a);
    WV_LAUNCH_CHECK();
  }
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
