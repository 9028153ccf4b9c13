// Tiled sm_100a correlation for the tracker's configuration: fp16 maps, C = 128, 3x3 patches, radius 3.
//
// Replaces corr_forward_kernel + the ATen blend (dpvo/altcorr/correlation_kernel.cu:82-136, 193-233) with
//
//   pass 1  nchw_to_nhwc_kernel: the frame-feature ring fmap2 [slot][C][H][W] is restaged channels-last
//           [slot][H][W][C] in the caller's workspace, so that one pixel's 128 channels are one aligned 256-B line
//           (in NCHW the contiguous unit is a ~20-B x-run per (c, y): every 32-B sector fetched is 2/3 waste).
//   pass 2  corr_tiled_kernel, persistent, one CTA per SM, warp-specialised:
//             producer warp   per edge: bounding box of the nine 8x8 tap windows -> two TMA tile loads
//                             (cp.async.bulk.tensor.4d, box {64 ch, BX, BY, 1}, 128-B swizzle, OOB zero fill ==
//                             the reference's within_bounds rule) into a 4-stage shared-memory ring, mbarrier
//                             complete_tx signalling.
//             8 consumer warps  per edge: D[9 x npix] = patch[9 x 128] . tile[128 x npix] with mma.sync m16n8k16
//                             (fp16 in, fp32 accumulate; B fragments by ldmatrix from the swizzled tile), then the
//                             4-corner bilinear blend of every patch pixel's own 8x8 sub-window and the permuted
//                             [x-tap][y-tap][3][3] fp16 store (correlation_kernel.cu:221-232), fused.
//
// Why tensor cores here (BASELINE north_star: "only if ncu shows that step is compute-bound"): the contraction is
// 2*128*64*9 = 147 kflop per edge and level, 13.4 GFLOP per two-level call; on the fp32 pipe (fp16 -> fp32 converts
// included) that is >= 0.4 ms on 148 SMs, above the 0.37 ms HBM time of the algorithmic bytes -- see profiles/.
// As a [16 x 128] x [128 x npix] MMA per edge it is ~1 % of the tensor pipe and the kernel goes back to being bound
// by L2 -> SM tile traffic.
//
// Edges whose windows do not fit the larger TMA box (patch magnified > ~2x between frames), and edges with an
// out-of-range index, take a per-warp direct-gather path inside the same kernel (same arithmetic, no tile).
#include <cuda.h>
#include <mma.h>

#include "wv_common.cuh"

namespace wv {

namespace {

constexpr int kC = 128, kP = 3, kPP = 9, kR = 3, kD = 8, kT = 7;
constexpr int kConsumers = 8;
constexpr int kThreads = 32 * (kConsumers + 1);
constexpr int kStages = 4;
constexpr int kBoxSmall = 10, kBoxLarge = 12;              // tile edge in pixels (two tensor maps)
constexpr int kMaxPix = kBoxLarge * kBoxLarge;             // 144
constexpr int kHalfBytes = kMaxPix * 128;                  // 18 432: one 64-channel half of a stage (1024-B multiple)
constexpr int kStageBytes = 2 * kHalfBytes;                // 36 864
constexpr int kRawStride = 152;                            // floats per patch-pixel row of D (bank-conflict-free stores)
constexpr int kRawBytes = kPP * kRawStride * 4;            // 5 472
constexpr int kPatchBytes = kC * kPP * 2;                  // 2 304
constexpr int kWarpBytes = kRawBytes + kPatchBytes;        // 7 776
constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kConsumers * kWarpBytes + (kConsumers + kStages) * 8 + 64;
constexpr int kClamp = 1 << 20;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  long long t0 = 0;
  int spins = 0;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    // watchdog: a protocol bug must surface as a CUDA error in seconds, not as a hung GPU (4e9 cycles ~ 2 s)
    if (!ok && (++spins & 0x3ff) == 0) {
      const long long now = clock64();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 4000000000LL) __trap();
    }
  } while (!ok);
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void mma_16816(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ int clamp_floor(float v) {
  // (int)floorf like the reference, clamped so that window arithmetic cannot overflow; NaN -> 0 like cvt.rzi
  const float f = floorf(v);
  if (!(f == f)) return 0;
  return (int)fminf(fmaxf(f, (float)-kClamp), (float)kClamp);
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// pass 1: [slot][C=128][HW] -> [slot][HW][C=128], 64-pixel tiles through shared memory
// ---------------------------------------------------------------------------------------------------
namespace {

constexpr int kTrPix = 64;
constexpr int kTrStrideW = 65;  // 32-bit words per pixel row in smem (64 + 1: odd stride)

__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const __half* __restrict__ src, __half* __restrict__ dst,
                                                           int HW, int tiles_per_slot, long long n_tiles, int vec_ok) {
  __shared__ uint32_t T[kTrPix * kTrStrideW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool even = vec_ok != 0;  // HW even and src 4-byte aligned: half2 loads
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long slot = tile / tiles_per_slot;
    const int p0 = (int)(tile % tiles_per_slot) * kTrPix;
    const __half* s = src + slot * (long long)kC * HW;
    // read: warp handles channel pairs (c, c+1); lane handles pixels p0 + 2*lane, +1
    for (int cp = warp; cp < kC / 2; cp += 8) {
      const int c = 2 * cp, px = p0 + 2 * lane;
      __half a0 = __float2half(0.f), a1 = a0, b0 = a0, b1 = a0;
      const __half* r0 = s + (long long)c * HW + px;
      const __half* r1 = r0 + HW;
      if (even && px + 1 < HW) {
        const __half2 v0 = *reinterpret_cast<const __half2*>(r0), v1 = *reinterpret_cast<const __half2*>(r1);
        a0 = __low2half(v0); a1 = __high2half(v0); b0 = __low2half(v1); b1 = __high2half(v1);
      } else {
        if (px < HW) { a0 = r0[0]; b0 = r1[0]; }
        if (px + 1 < HW) { a1 = r0[1]; b1 = r1[1]; }
      }
      const __half2 w0 = __halves2half2(a0, b0), w1 = __halves2half2(a1, b1);  // (c, c+1) of pixel px / px+1
      T[(2 * lane) * kTrStrideW + cp] = *reinterpret_cast<const uint32_t*>(&w0);
      T[(2 * lane + 1) * kTrStrideW + cp] = *reinterpret_cast<const uint32_t*>(&w1);
    }
    __syncthreads();
    // write: one pixel = 64 words = 256 B; a warp writes 2 x 128 B per pixel
    uint32_t* d = reinterpret_cast<uint32_t*>(dst + (slot * (long long)HW + p0) * kC);
    for (int t = threadIdx.x; t < kTrPix * 64; t += 256) {
      const int px = t >> 6, wd = t & 63;
      if (p0 + px < HW) d[t] = T[px * kTrStrideW + wd];
    }
    __syncthreads();
  }
}

struct CorrParams {
  const __half* fmap1;  // [B][n1][128][3][3]
  const __half* fmap2;  // [B][n2][128][H][W]  (original layout, direct-gather path only)
  const float* coords;  // [B][E][2][3][3]
  const int64_t* ii;
  const int64_t* jj;
  __half* out;          // [B][E][7][7][3][3]
  int B, E, n1, n2, H, W;
  int* status;
};

// geometry of one edge, evaluated redundantly by the producer warp and by the consumer warp that owns the edge
struct EdgeGeom {
  int fx, fy;      // lane p < 9: integer tap origin of patch pixel p (floor of its coordinate)
  float dx, dy;    // lane p < 9: bilinear fractions
  int x0, y0;      // warp-uniform: top-left of the tile (min floor - R)
  int kind;        // warp-uniform: 0 small box, 1 large box, 2 direct gather, 3 nothing to read (all taps outside / bad index)
  int slot;        // warp-uniform: b * n2 + jj
  long long patch; // warp-uniform: b * n1 + ii
};

// the raw global loads of one edge, issued one iteration ahead of their use (software prefetch)
struct EdgeRaw { float cv; int64_t ix, jx; long long be; };

__device__ __forceinline__ EdgeRaw edge_load(const CorrParams& P, long long be, int lane) {
  EdgeRaw r;
  const int e = (int)(be % P.E);
  r.be = be;
  r.ix = P.ii[e]; r.jx = P.jj[e];
  r.cv = lane < 2 * kPP ? P.coords[be * (2 * kPP) + lane] : 0.f;
  return r;
}

__device__ __forceinline__ EdgeGeom edge_geom(const CorrParams& P, const EdgeRaw& raw, int lane) {
  EdgeGeom g;
  const int b = (int)(raw.be / P.E);
  const int64_t ix = raw.ix, jx = raw.jx;
  const float cv = raw.cv;
  const float x = __shfl_sync(0xffffffffu, cv, lane < kPP ? lane : 0);
  const float y = __shfl_sync(0xffffffffu, cv, lane < kPP ? lane + kPP : kPP);
  g.fx = clamp_floor(x); g.fy = clamp_floor(y);
  g.dx = x - floorf(x); g.dy = y - floorf(y);
  int mnx = lane < kPP ? g.fx : kClamp * 2, mxx = lane < kPP ? g.fx : -kClamp * 2;
  int mny = lane < kPP ? g.fy : kClamp * 2, mxy = lane < kPP ? g.fy : -kClamp * 2;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) {
    mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
    mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, o)); mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
  }
  mnx = __shfl_sync(0xffffffffu, mnx, 0); mxx = __shfl_sync(0xffffffffu, mxx, 0);
  mny = __shfl_sync(0xffffffffu, mny, 0); mxy = __shfl_sync(0xffffffffu, mxy, 0);
  g.x0 = mnx - kR; g.y0 = mny - kR;
  const int wdt = mxx - mnx + kD, hgt = mxy - mny + kD;
  const bool bad = ix < 0 || ix >= P.n1 || jx < 0 || jx >= P.n2;
  const bool outside = g.x0 >= P.W || g.y0 >= P.H || g.x0 + wdt <= 0 || g.y0 + hgt <= 0;
  if (bad || outside) g.kind = 3;
  else if (wdt <= kBoxSmall && hgt <= kBoxSmall) g.kind = 0;
  else if (wdt <= kBoxLarge && hgt <= kBoxLarge) g.kind = 1;
  else g.kind = 2;
  if (bad && lane == 0) set_status(P.status, WV_STATUS_BAD_INDEX);
  g.slot = bad ? 0 : b * P.n2 + (int)jx;
  g.patch = bad ? 0 : (long long)b * P.n1 + ix;
  return g;
}

// the producer's view of an edge, computed by ONE lane (32 edges per warp pass, loads in flight together)
__device__ __forceinline__ void edge_box(const CorrParams& P, long long be, int& kind, int& x0, int& y0, int& slot) {
  const int b = (int)(be / P.E), e = (int)(be % P.E);
  const int64_t ix = P.ii[e], jx = P.jj[e];
  const float2* c2 = reinterpret_cast<const float2*>(P.coords + be * (2 * kPP));  // 72-B records: 8-byte aligned
  float v[2 * kPP];
#pragma unroll
  for (int i = 0; i < kPP; i++) { const float2 t = c2[i]; v[2 * i] = t.x; v[2 * i + 1] = t.y; }
  int mnx = kClamp * 2, mxx = -kClamp * 2, mny = kClamp * 2, mxy = -kClamp * 2;
#pragma unroll
  for (int p = 0; p < kPP; p++) {
    const int fx = clamp_floor(v[p]), fy = clamp_floor(v[kPP + p]);
    mnx = min(mnx, fx); mxx = max(mxx, fx); mny = min(mny, fy); mxy = max(mxy, fy);
  }
  x0 = mnx - kR; y0 = mny - kR;
  const int wdt = mxx - mnx + kD, hgt = mxy - mny + kD;
  const bool bad = ix < 0 || ix >= P.n1 || jx < 0 || jx >= P.n2;
  const bool outside = x0 >= P.W || y0 >= P.H || x0 + wdt <= 0 || y0 + hgt <= 0;
  if (bad || outside) kind = 3;
  else if (wdt <= kBoxSmall && hgt <= kBoxSmall) kind = 0;
  else if (wdt <= kBoxLarge && hgt <= kBoxLarge) kind = 1;
  else kind = 2;
  slot = bad ? 0 : b * P.n2 + (int)jx;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// pass 2: persistent warp-specialised kernel
// ---------------------------------------------------------------------------------------------------
namespace {

// NT (<= 4) pixel tiles of 8 px: D[16 x 8*NT] = A[16 x 128] . tile^T, B fragments double-buffered one 32-channel step ahead.
// a0 / a1: this lane's ldmatrix address in tile 0 for chunks 0..3 / 4..7 of the 128-B rows (swizzle folded in).
template <int NT>
__device__ __forceinline__ void mma_group(uint32_t a0, uint32_t a1, const uint32_t (&A)[8][4], float* d0, float* d8, bool has8) {
  float acc[NT][4];
  uint32_t b[2][NT][4];
#pragma unroll
  for (int u = 0; u < NT; u++) {
    acc[u][0] = 0.f; acc[u][1] = 0.f; acc[u][2] = 0.f; acc[u][3] = 0.f;
    ldmatrix_x4(a0 + u * 1024, b[0][u][0], b[0][u][1], b[0][u][2], b[0][u][3]);
  }
#pragma unroll
  for (int hq = 0; hq < 4; hq++) {  // channels 32*hq .. 32*hq+31 = k-steps 2*hq, 2*hq+1
    if (hq < 3) {
      const uint32_t an = (((hq + 1) & 1) ? a1 : a0) + ((hq + 1) >> 1) * kHalfBytes;
#pragma unroll
      for (int u = 0; u < NT; u++)
        ldmatrix_x4(an + u * 1024, b[(hq + 1) & 1][u][0], b[(hq + 1) & 1][u][1], b[(hq + 1) & 1][u][2], b[(hq + 1) & 1][u][3]);
    }
#pragma unroll
    for (int u = 0; u < NT; u++) mma_16816(acc[u], A[2 * hq], b[hq & 1][u][0], b[hq & 1][u][1]);
#pragma unroll
    for (int u = 0; u < NT; u++) mma_16816(acc[u], A[2 * hq + 1], b[hq & 1][u][2], b[hq & 1][u][3]);
  }
#pragma unroll
  for (int u = 0; u < NT; u++) {
    *reinterpret_cast<float2*>(d0 + 8 * u) = make_float2(acc[u][0], acc[u][1]);
    if (has8) *reinterpret_cast<float2*>(d8 + 8 * u) = make_float2(acc[u][2], acc[u][3]);
  }
}

__global__ void __launch_bounds__(kThreads, 1)
corr_tiled_kernel(const __grid_constant__ CUtensorMap map_small, const __grid_constant__ CUtensorMap map_large, const CorrParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // stays a shared-space pointer (LDS/STS, not generic LD/ST)
  uint8_t* stages = smem;                                              // [kStages][2][kHalfBytes], 1024-B aligned
  uint8_t* warp_area = stages + kStages * kStageBytes;                 // [kConsumers][raw | patch]
  // Barriers.  full[w]: one per CONSUMER WARP (not per stage) -- tile q goes to warp q % 8 and stage q % 4, so a warp's
  // consecutive tiles reuse one stage and its barrier can never run a phase ahead of the warp; a per-stage barrier would be
  // waited on by two warps that are a phase apart, which a parity wait cannot tell from "already complete".
  // empty[s]: one per stage, waited on by the producer alone, in order.
  uint64_t* bars = reinterpret_cast<uint64_t*>(warp_area + kConsumers * kWarpBytes);  // full[kConsumers], empty[kStages]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + kConsumers);
  static_assert(kConsumers % kStages == 0, "a warp's tiles must always land in the same stage");

  if (threadIdx.x == 0) {
    for (int w = 0; w < kConsumers; w++) mbar_init(full0 + 8 * w, 1);
    for (int s = 0; s < kStages; s++) mbar_init(empty0 + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  const long long BE = (long long)P.B * P.E;
  const long long lo = BE * blockIdx.x / gridDim.x, hi = BE * (blockIdx.x + 1) / gridDim.x;
  const int nt = (int)(hi - lo);
  const int per = (nt + kConsumers - 1) / kConsumers;  // edges per consumer warp (contiguous runs: patch reuse)

  if (warp == kConsumers) {
    // ================================================================= producer
    const int total = per * kConsumers;
    for (int q0 = 0; q0 < total; q0 += 32) {
      // every lane works out the box of one tile of this batch (their global loads overlap) ...
      int kind = 3, x0 = 0, y0 = 0, slot = 0;
      {
        const int q = q0 + lane, idx = (q % kConsumers) * per + q / kConsumers;
        if (q < total && idx < nt) edge_box(P, lo + idx, kind, x0, y0, slot);
      }
      // ... then lane 0 feeds the ring in order
      const int nb = min(32, total - q0);
      for (int i = 0; i < nb; i++) {
        const int q = q0 + i, w = q % kConsumers, s = q % kStages;
        const int kd = __shfl_sync(0xffffffffu, kind, i), bx = __shfl_sync(0xffffffffu, x0, i);
        const int by = __shfl_sync(0xffffffffu, y0, i), sl = __shfl_sync(0xffffffffu, slot, i);
        if (lane == 0) {
          mbar_wait(empty0 + 8 * s, ((q / kStages) & 1) ^ 1);
          const uint32_t bar = full0 + 8 * w;
          if (kd <= 1) {
            const int box = kd == 0 ? kBoxSmall : kBoxLarge;
            const CUtensorMap* map = kd == 0 ? &map_small : &map_large;
            const uint32_t dst = smem_u32(stages + (size_t)s * kStageBytes);
            mbar_arrive_expect_tx(bar, (uint32_t)(2 * box * box * 128));
            tma_load_4d(dst, map, bar, 0, bx, by, sl);
            tma_load_4d(dst + kHalfBytes, map, bar, 64, bx, by, sl);
          } else {
            mbar_arrive(bar);
          }
        }
      }
      __syncwarp();
    }
    return;
  }

  // =================================================================== consumers
  float* rawD = reinterpret_cast<float*>(warp_area + (size_t)warp * kWarpBytes);  // [9][kRawStride]: cols 0..143 D, 144..146 geometry
  __half* patch = reinterpret_cast<__half*>(warp_area + (size_t)warp * kWarpBytes + kRawBytes);  // [128][9]
  const int g8 = lane >> 2, t4 = lane & 3;
  const uint32_t lds_off0 = (uint32_t)((lane & 7) * 128 + ((((lane >> 3) ^ (lane & 7)) & 7) << 4));  // ldmatrix row + swizzled chunk
  const uint32_t lds_off1 = lds_off0 ^ 64u;                                                            // chunks 4..7 of the 128-B row
  float* const d0 = rawD + g8 * kRawStride + 2 * t4;
  float* const d8 = rawD + 8 * kRawStride + 2 * t4;
  uint32_t A[8][4];
  long long cur_patch = -1, pf_patch = -1;
  uint4 pf[5];
  const int cnt = max(0, min(per, nt - warp * per));
  EdgeRaw raw{};
  if (cnt > 0) raw = edge_load(P, lo + warp * per, lane);

  for (int k = 0; k < per; k++) {
    const int q = k * kConsumers + warp, s = q % kStages;
    const uint32_t fullb = full0 + 8 * warp, emptyb = empty0 + 8 * s;
    if (k >= cnt) {  // padding slot of the ring: keep the protocol in step
      mbar_wait(fullb, k & 1);
      if (lane == 0) mbar_arrive(emptyb);
      continue;
    }
    const long long be = lo + warp * per + k;
    const EdgeGeom g = edge_geom(P, raw, lane);
    if (k + 1 < cnt) raw = edge_load(P, be + 1, lane);  // next edge's loads fly during this edge's work

    // ---- A operand: the 3x3x128 source patch (reused while consecutive edges share it; prefetched otherwise)
    if (g.kind != 3 && g.patch != cur_patch) {
      uint4* dstp = reinterpret_cast<uint4*>(patch);
      if (pf_patch != g.patch) {
        const uint4* src = reinterpret_cast<const uint4*>(P.fmap1 + g.patch * (kC * kPP));
#pragma unroll
        for (int i = 0; i < 5; i++) if (lane + 32 * i < kPatchBytes / 16) pf[i] = src[lane + 32 * i];
      }
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 5; i++) if (lane + 32 * i < kPatchBytes / 16) dstp[lane + 32 * i] = pf[i];
      __syncwarp();
      const unsigned short* ps = reinterpret_cast<const unsigned short*>(patch);
#pragma unroll
      for (int ks = 0; ks < 8; ks++) {
        const int c0 = 16 * ks + 2 * t4;
        A[ks][0] = (uint32_t)ps[c0 * kPP + g8] | ((uint32_t)ps[(c0 + 1) * kPP + g8] << 16);
        A[ks][2] = (uint32_t)ps[(c0 + 8) * kPP + g8] | ((uint32_t)ps[(c0 + 9) * kPP + g8] << 16);
        A[ks][1] = 0u; A[ks][3] = 0u;
        if (g8 == 0) {  // row 8 of the 16-row fragment = patch pixel 8; rows 9..15 stay zero
          A[ks][1] = (uint32_t)ps[c0 * kPP + 8] | ((uint32_t)ps[(c0 + 1) * kPP + 8] << 16);
          A[ks][3] = (uint32_t)ps[(c0 + 8) * kPP + 8] | ((uint32_t)ps[(c0 + 9) * kPP + 8] << 16);
        }
      }
      cur_patch = g.patch;
    }

    // ---- geometry of each patch pixel next to its D row
    int BX = kD, wx = 0, wy = 0;
    if (g.kind <= 1) { BX = g.kind == 0 ? kBoxSmall : kBoxLarge; wx = g.fx - kR - g.x0; wy = g.fy - kR - g.y0; }
    if (lane < kPP) {
      float* gr = rawD + lane * kRawStride + kMaxPix;
      gr[0] = __int_as_float(wy * BX + wx); gr[1] = g.dx; gr[2] = g.dy;
    }

    mbar_wait(fullb, k & 1);
    if (g.kind <= 1) {
      // -------------------------------------------------------------- tensor-core contraction from the TMA tile
      const uint32_t st = smem_u32(stages + (size_t)s * kStageBytes);
      const uint32_t a0 = st + lds_off0, a1 = st + lds_off1;
      const bool has8 = g8 == 0;
      mma_group<4>(a0, a1, A, d0, d8, has8);
      mma_group<4>(a0 + 4 * 1024, a1 + 4 * 1024, A, d0 + 32, d8 + 32, has8);
      mma_group<4>(a0 + 8 * 1024, a1 + 8 * 1024, A, d0 + 64, d8 + 64, has8);
      if (g.kind == 0) {
        mma_group<1>(a0 + 12 * 1024, a1 + 12 * 1024, A, d0 + 96, d8 + 96, has8);   // 13 tiles: 100 px (+4 unused)
      } else {
        mma_group<4>(a0 + 12 * 1024, a1 + 12 * 1024, A, d0 + 96, d8 + 96, has8);
        mma_group<2>(a0 + 16 * 1024, a1 + 16 * 1024, A, d0 + 128, d8 + 128, has8);  // 18 tiles: 144 px
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(emptyb);
    } else {
      if (lane == 0) mbar_arrive(emptyb);
      if (g.kind == 2) {
        // ------------------------------------------------------------ direct gather (windows too far apart for a tile)
        const size_t plane = (size_t)P.H * P.W;
        const __half* f2 = P.fmap2 + (size_t)g.slot * kC * plane;
        for (int t = lane; t < kPP * kD * kD; t += 32) {
          const int b = t & 7, a = (t >> 3) & 7, p = t >> 6;
          const int i1 = __shfl_sync(0xffffffffu, g.fy, p) + a - kR, j1 = __shfl_sync(0xffffffffu, g.fx, p) + b - kR;
          float sacc = 0.f;
          if (i1 >= 0 && i1 < P.H && j1 >= 0 && j1 < P.W) {
            const __half* p2 = f2 + (size_t)i1 * P.W + j1;
            for (int c = 0; c < kC; c++) sacc = fmaf(__half2float(patch[c * kPP + p]), __half2float(p2[(size_t)c * plane]), sacc);
          }
          rawD[p * kRawStride + a * kD + b] = sacc;
        }
      } else {
        for (int t = lane; t < kPP * kD * kD; t += 32) rawD[(t >> 6) * kRawStride + (t & 63)] = 0.f;
      }
      __syncwarp();
    }

    // ---- prefetch the next edge's patch into registers; the blend below hides its latency
    if (k + 1 < cnt) {
      const int64_t nix = raw.ix;
      const long long np = (long long)((be + 1) / P.E) * P.n1 + nix;
      if (nix >= 0 && nix < P.n1 && np != cur_patch && np != pf_patch) {
        const uint4* src = reinterpret_cast<const uint4*>(P.fmap1 + np * (kC * kPP));
#pragma unroll
        for (int i = 0; i < 5; i++) if (lane + 32 * i < kPatchBytes / 16) pf[i] = src[lane + 32 * i];
        pf_patch = np;
      }
    }

    // ---- bilinear blend of each patch pixel's own 8x8 window + permuted store (correlation_kernel.cu:221-232).
    // item = (patch pixel p, x-tap b): 63 items, two per lane; separable: 8 horizontal lerps, then 7 vertical ones
    __half* o = P.out + be * (kT * kT * kPP);
#pragma unroll
    for (int half_i = 0; half_i < 2; half_i++) {
      const int item = lane + 32 * half_i;
      if (item < kPP * kT) {
        const int p = item % kPP, b = item / kPP;
        const float* row = rawD + p * kRawStride;
        const int org = __float_as_int(row[kMaxPix]);
        const float dx = row[kMaxPix + 1], dy = row[kMaxPix + 2];
        const float* rr = row + org + b;
        float Hh[kD];
#pragma unroll
        for (int a = 0; a < kD; a++) {
          const float r0 = rr[0], r1 = rr[1];
          Hh[a] = fmaf(dx, r1 - r0, r0);
          rr += BX;
        }
        __half* op = o + b * (kT * kPP) + p;
#pragma unroll
        for (int a = 0; a < kT; a++) op[a * kPP] = __float2half_rn(fmaf(dy, Hh[a + 1] - Hh[a], Hh[a]));
      }
    }
    __syncwarp();
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
    else
      (void)cudaGetLastError();
  }
  return fn;
}

bool make_map(CUtensorMap* m, void* nhwc, int W, int H, long long slots, int box) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  const cuuint64_t dims[4] = {(cuuint64_t)kC, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)slots};
  const cuuint64_t strides[3] = {(cuuint64_t)kC * 2, (cuuint64_t)kC * 2 * W, (cuuint64_t)kC * 2 * W * H};
  const cuuint32_t boxd[4] = {64, (cuuint32_t)box, (cuuint32_t)box, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, nhwc, dims, strides, boxd, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int tiled_mode() {  // WVB200_CORR_TILED: unset = heuristic, 0 = never, 1 = whenever the shape allows
  const char* s = getenv("WVB200_CORR_TILED");
  if (!s || !s[0]) return -1;
  return s[0] == '0' ? 0 : 1;
}

}  // namespace

bool corr_tiled_supported(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int R) {
  const int mode = tiled_mode();
  if (mode == 0) return false;
  if (dtype != WV_F16 || C != kC || PH != kP || PW != kP || R != kR) return false;
  if (B <= 0 || E <= 0 || n1 <= 0 || (long long)B * n2 > 0x7fffffffLL || H2 > (1 << 20) || W2 > (1 << 20)) return false;
  if (!encode_tiled_fn()) return false;
  if (mode == 1) return true;
  // restaging the ring costs one read + one write of it: worth it once the edges' tile traffic dominates
  const double ring = (double)B * n2 * H2 * W2 * kC * 2.0, tiles = (double)B * E * 2.0 * kBoxSmall * kBoxSmall * 128.0;
  return (long long)B * E >= 1024 && tiles >= 0.5 * ring;
}

size_t corr_tiled_workspace_bytes(int, int B, int, int, int, int, int, int n2, int H2, int W2, int) {
  return (size_t)B * n2 * H2 * W2 * kC * 2 + 2048;
}

int corr_forward_tiled(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii, const int64_t* jj, void* out,
                       int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int R, void* workspace,
                       size_t workspace_bytes, int* status, cudaStream_t st) {
  if (!corr_tiled_supported(dtype, B, E, C, PH, PW, n1, n2, H2, W2, R)) return WV_EUNSUPPORTED;
  if (workspace_bytes < corr_tiled_workspace_bytes(dtype, B, E, C, PH, PW, n1, n2, H2, W2, R)) return WV_EWORKSPACE;
  if ((reinterpret_cast<uintptr_t>(fmap1) & 15) || (reinterpret_cast<uintptr_t>(fmap2) & 1)) return WV_EUNSUPPORTED;
  __half* nhwc = reinterpret_cast<__half*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~(uintptr_t)1023);
  const long long slots = (long long)B * n2;
  const int HW = H2 * W2;
  CUtensorMap ms, ml;
  if (!make_map(&ms, nhwc, W2, H2, slots, kBoxSmall) || !make_map(&ml, nhwc, W2, H2, slots, kBoxLarge)) return WV_EUNSUPPORTED;

  const int tiles_per_slot = (HW + kTrPix - 1) / kTrPix;
  const long long n_tiles = slots * tiles_per_slot;
  const long long cap = (long long)sm_count() * 8;
  const int vec_ok = ((HW & 1) == 0 && (reinterpret_cast<uintptr_t>(fmap2) & 3) == 0) ? 1 : 0;
  nchw_to_nhwc_kernel<<<(int)(n_tiles < cap ? n_tiles : cap), 256, 0, st>>>((const __half*)fmap2, nhwc, HW, tiles_per_slot, n_tiles, vec_ok);
  WV_LAUNCH_CHECK();

  static bool attr_set = false;
  if (!attr_set) {
    WV_CUDA_TRY(cudaFuncSetAttribute(corr_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  CorrParams P{(const __half*)fmap1, (const __half*)fmap2, coords, ii, jj, (__half*)out, B, E, n1, n2, H2, W2, status};
  const long long BE = (long long)B * E;
  const long long want = (BE + kConsumers - 1) / kConsumers;
  const int grid = (int)(want < sm_count() ? want : sm_count());
  corr_tiled_kernel<<<grid, kThreads, kSmemBytes, st>>>(ms, ml, P);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

}  // namespace wv
