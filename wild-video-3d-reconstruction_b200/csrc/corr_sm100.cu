// Tiled sm_100a correlation for the tracker's configuration: fp16 maps, C = 128, 3x3 patches, radius 3.
//
// Replaces corr_forward_kernel + the ATen blend (dpvo/altcorr/correlation_kernel.cu:82-136, 193-233) with
//
//   pass 1  restage_kernel: the frame-feature ring fmap2 [slot][C][H][W] is restaged channels-last [slot][H][W][C] in the
//           caller's workspace, so that one pixel's 128 channels are one aligned 256-B line (in NCHW the contiguous unit
//           is a ~20-B x-run per (c, y): every 32-B sector fetched is 2/3 waste); the source patches gmap [n][C][3][3] are
//           restaged pixel-major: [n][2 halves][9][64] rows of 128 B with the 128-B swizzle applied (the tcgen05 B operand,
//           two bulk copies), or plain [n][9][C] for the mma.sync variant.
//   pass 2  corr_tc5_kernel (default), persistent, one CTA per SM, three independent pipelines of six warps each:
//             producer warp   owns all global reads.  16 lanes prepare one edge each: bounding box of the nine 8x8 tap
//                             windows, per-pixel window origin + bilinear fractions -> a 104-B record.  Then, per edge,
//                             into the next of the pipeline's 2 stages: two TMA tile loads (cp.async.bulk.tensor.4d,
//                             64 channels x 10x10 px, 128-B swizzle, OOB zero fill == the reference's within_bounds rule)
//                             and two bulk copies of the pre-swizzled patch halves, all completing on the stage's
//                             mbarrier.  Windows up to 12x12 px go as two 12x6 half tiles (two slots).
//             MMA warp        one lane: D[128 px x 16] = tile[128 px x 128 ch] . patch[16 x 128 ch]^T as 8 tcgen05.mma
//                             (M128 N16 K16, fp16 in, fp32 accumulate in TMEM); the TMA-written tile is the K-major A
//                             operand in place; tcgen05.commit releases the stage and publishes the accumulator.
//             4 consumer warps  tcgen05.ld of their TMEM lane quarter, the 9 real columns scattered to a per-edge fp32
//                             scratch, then the 4-corner bilinear blend of every patch pixel's own 8x8 sub-window and
//                             the permuted [x-tap][y-tap][3][3] fp16 store (correlation_kernel.cu:221-232), fused.
//           Tiles are 9x9 px when the window fits (always at the coarse level), else 10x10, else 12x12 in two halves.
//           corr_tc5_kernel<2> is the two-level form (DPVO.corr, dpvo/dpvo.py:325-333): level 0 and level 1 of an edge go through
//           the pipeline back to back and the (level 0, level 1) pairs are stored once in the torch.stack layout.
//           corr_tiled_kernel (WVB200_CORR_TC5=0) is the earlier form of pass 2 with the contraction on mma.sync m16n8k16
//           inside 8 consumer warps; it is kept as a cross-check of the tcgen05 path (tests run both).
//
// Why tensor cores here (BASELINE north_star: "only if ncu shows that step is compute-bound"): the contraction is
// 2*128*64*9 = 147 kflop per edge and level, 13.4 GFLOP per two-level call; on the fp32 pipe (fp16 -> fp32 converts
// included) that is >= 0.4 ms on 148 SMs, above the 0.37 ms HBM time of the algorithmic bytes -- see profiles/.
// mma.sync measured ~66 us of tensor-pipe time per level on the warps that also blend (104 warp-wide MMAs per edge, 7/16
// of the A rows padding); with tcgen05 the MMAs cost one thread 8 instructions per edge and the kernel is bound by
// L2 -> SM tile traffic and the blend.
//
// Edges whose windows do not fit 12x12 px (patch magnified > ~2x between frames) take a direct-gather path
// inside the same kernel (same arithmetic, no tile); edges with an out-of-range index produce zeros and set the status.
#include <cuda.h>
#include <mma.h>

#include "wv_common.cuh"

namespace wv {

namespace {

constexpr int kC = 128, kP = 3, kPP = 9, kR = 3, kD = 8, kT = 7;
constexpr int kConsumers = 8;
constexpr int kProducers = 2;                              // each feeds kConsumers / kProducers consumer warps through its own stages
constexpr int kThreads = 32 * (kConsumers + kProducers);
constexpr int kStages = 6;
constexpr int kBoxSmall = 10, kBoxLarge = 12, kBoxLargeRows = 6;  // tile shapes: 10x10, or 12x12 as two 12x6 halves
constexpr int kMaxPix = kBoxLarge * kBoxLarge;             // 144 columns of D at most
constexpr int kHalfBytes = 13 * 1024;                      // one 64-channel half of a stage: 100 px x 128 B = 12 800 (1024-B aligned slots)
constexpr int kPatchOff = 2 * kHalfBytes;                  // 26 624: the [9][128] fp16 patch
constexpr int kPatchBytes = kC * kPP * 2;                  // 2 304
constexpr int kStageBytes = 29 * 1024;                     // 29 696
constexpr int kRawStride = 152;                            // floats per patch-pixel row of D (bank-conflict-free stores)
constexpr int kRawBytes = kPP * kRawStride * 4;            // 5 472
constexpr int kRecBytes = 96;                              // per-edge record (below)
constexpr int kRecRing = 2 * 32 * kRecBytes;               // per producer: two batches of 16 records
constexpr int kBarBytes = (kStages + 2 * kConsumers) * 8 + kConsumers * 4 + 32;
constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kConsumers * kRawBytes + kRecRing + kBarBytes;
constexpr int kClamp = 1 << 20;
static_assert(kBoxSmall * kBoxSmall * 128 <= kHalfBytes && kBoxLarge * kBoxLargeRows * 128 <= kHalfBytes, "tile must fit its slot");
static_assert(kPatchOff + kPatchBytes <= kStageBytes, "patch must fit its slot");
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");

// the record the producer leaves for the consumer of an edge (all 32 lanes write theirs at once)
struct EdgeRec {
  int flags;                  // kind (bits 0-1): 0 small tile, 1 large tile in two halves, 2 direct gather, 3 nothing to read; bit 2: load the patch
  unsigned char org[12];      // column of pixel p's window origin in the D column space (< 144)
  float dx[kPP], dy[kPP];     // bilinear fractions
};
static_assert(sizeof(EdgeRec) <= kRecBytes, "record slot");

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
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"   // %3: suspend-time hint (ns): fewer polls compete for issue slots
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(4000u)
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
// pass 1: restaging.  Blocks [0, n_tiles): fmap2 [slot][128][HW] -> [slot][HW][128] in 64-pixel tiles through shared
// memory; blocks beyond: gmap [n][128][9] -> [n][9][128], one warp per patch.
// ---------------------------------------------------------------------------------------------------
namespace {

constexpr int kTrPix = 64;
constexpr int kTrStrideW = 65;  // 32-bit words per pixel row in smem (64 + 1: odd stride)

__global__ void __launch_bounds__(256) restage_kernel(const __half* __restrict__ fmap2, __half* __restrict__ nhwc, int HW,
                                                      int tiles_per_slot, long long n_tiles, int vec_ok,
                                                      const __half* __restrict__ gmap, __half* __restrict__ gT, long long n_patches,
                                                      int tile_blocks, int patch_kmajor) {
  __shared__ uint32_t T[kTrPix * kTrStrideW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if ((int)blockIdx.x >= tile_blocks) {
    // ---- patches: out word (p, channel pair) = (in[2cp][p], in[2cp+1][p])
    const long long w0 = (long long)(blockIdx.x - tile_blocks) * 8 + warp, wstep = (long long)(gridDim.x - tile_blocks) * 8;
    for (long long n = w0; n < n_patches; n += wstep) {
      const unsigned short* in = reinterpret_cast<const unsigned short*>(gmap) + n * (kC * kPP);
      uint32_t* out = reinterpret_cast<uint32_t*>(gT + n * (kC * kPP));
      if (patch_kmajor) {
        // tcgen05 B operand: [64-channel half][patch pixel p][64 ch] = rows of 128 B with the 128-B swizzle already applied, so
        // that two plain bulk copies (9 rows x 128 B each) land it in canonical K-major form
#pragma unroll
        for (int i = 0; i < kPP * 64 / 32; i++) {
          const int o = i * 32 + lane, h = o / (kPP * 32), r = o - h * (kPP * 32), p = r >> 5, w = r & 31;
          const int cp = h * 32 + (((w >> 2) ^ (p & 7)) & 7) * 4 + (w & 3);
          out[o] = (uint32_t)in[(2 * cp) * kPP + p] | ((uint32_t)in[(2 * cp + 1) * kPP + p] << 16);
        }
      } else {
#pragma unroll
        for (int i = 0; i < kPP * 64 / 32; i++) {
          const int o = i * 32 + lane, p = o >> 6, cp = o & 63;
          out[o] = (uint32_t)in[(2 * cp) * kPP + p] | ((uint32_t)in[(2 * cp + 1) * kPP + p] << 16);
        }
      }
    }
    return;
  }
  for (long long tile = blockIdx.x; tile < n_tiles; tile += tile_blocks) {
    const long long slot = tile / tiles_per_slot;
    const int p0 = (int)(tile % tiles_per_slot) * kTrPix;
    const __half* s = fmap2 + slot * (long long)kC * HW;
    if (vec_ok && p0 + kTrPix <= HW) {
      // 16-B loads: work item = (channel pair cp, 8-pixel group xg); a warp covers 4 channel rows x 128 B per load
#pragma unroll
      for (int it = 0; it < 2; it++) {
        const int xg = lane & 7, cp = it * 32 + warp * 4 + (lane >> 3);
        const uint4 a = *reinterpret_cast<const uint4*>(s + (long long)(2 * cp) * HW + p0 + 8 * xg);
        const uint4 b = *reinterpret_cast<const uint4*>(s + (long long)(2 * cp + 1) * HW + p0 + 8 * xg);
        uint32_t* t = T + (8 * xg) * kTrStrideW + cp;
        t[0 * kTrStrideW] = __byte_perm(a.x, b.x, 0x5410); t[1 * kTrStrideW] = __byte_perm(a.x, b.x, 0x7632);
        t[2 * kTrStrideW] = __byte_perm(a.y, b.y, 0x5410); t[3 * kTrStrideW] = __byte_perm(a.y, b.y, 0x7632);
        t[4 * kTrStrideW] = __byte_perm(a.z, b.z, 0x5410); t[5 * kTrStrideW] = __byte_perm(a.z, b.z, 0x7632);
        t[6 * kTrStrideW] = __byte_perm(a.w, b.w, 0x5410); t[7 * kTrStrideW] = __byte_perm(a.w, b.w, 0x7632);
      }
    } else {
      // ragged / unaligned maps: scalar loads
      const unsigned short* su = reinterpret_cast<const unsigned short*>(s);
      for (int cp = warp; cp < kC / 2; cp += 8) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int px = p0 + lane + 32 * h;
          uint32_t v = 0;
          if (px < HW) v = (uint32_t)su[(long long)(2 * cp) * HW + px] | ((uint32_t)su[(long long)(2 * cp + 1) * HW + px] << 16);
          T[(lane + 32 * h) * kTrStrideW + cp] = v;
        }
      }
    }
    __syncthreads();
    // write: one pixel = 64 words = 256 B; a warp writes 128 B per instruction
    uint32_t* d = reinterpret_cast<uint32_t*>(nhwc + (slot * (long long)HW + p0) * kC);
#pragma unroll
    for (int it = 0; it < 16; it++) {
      const int t = it * 256 + threadIdx.x, px = t >> 6, wd = t & 63;
      if (p0 + px < HW) d[t] = T[px * kTrStrideW + wd];
    }
    __syncthreads();
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// pass 2: persistent warp-specialised kernel
// ---------------------------------------------------------------------------------------------------
namespace {

struct CorrParams {
  const __half* gT;     // [B][n1][9][128]   restaged patches (workspace)
  const __half* fmap2;  // [B][n2][128][H][W] original layout (direct-gather path only)
  const float* coords;  // [B][E][2][3][3]
  const int64_t* ii;
  const int64_t* jj;
  __half* out;          // [B][E][7][7][3][3]; two-level form: [B][E][7][7][3][3][2]
  int B, E, n1, n2, H, W;
  int* status;
  // second pyramid level of the two-level (stacked) form: its own map ring, extent and coordinate multiplier
  const __half* fmap2_l1;
  int H1, W1;
  float cmul0, cmul1;
};

struct CorrMaps {  // [level][10x10 | 12x6 half | 9x9]
  CUtensorMap m[6];
};
constexpr int kBoxTiny = 9;   // windows whose patch pixels fall within one pixel of each other (the coarse level): 81 instead of 100 rows

__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// NT pixel tiles of 8 px: D[16 x 8*NT] = A[16 x 128] . tile^T, B fragments double-buffered one 32-channel step ahead.
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
corr_tiled_kernel(const __grid_constant__ CUtensorMap map_small, const __grid_constant__ CUtensorMap map_half, const CorrParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // stays a shared-space pointer
  uint8_t* stages = smem;                                              // [kStages][kStageBytes], 1024-B aligned
  uint8_t* raw_area = stages + kStages * kStageBytes;                  // [kConsumers][kRawBytes]
  uint64_t* bars = reinterpret_cast<uint64_t*>(raw_area + kConsumers * kRawBytes);
  // Barriers.  full[w] (per CONSUMER WARP): the tile, patch and record of warp w's next slot have landed.  taken[w]: warp w
  // has picked that slot up -- the producer keeps at most one slot per warp outstanding, so a warp never waits on a
  // barrier that is a phase ahead of it (parity waits cannot tell "two phases ago" from "now").  empty[s] (per stage): the
  // consumer is done reading stage s; waited on by the producer alone, in ring order.
  const uint32_t full0 = smem_u32(bars), taken0 = full0 + 8 * kConsumers, empty0 = taken0 + 8 * kConsumers;
  volatile int* mailbox = reinterpret_cast<volatile int*>(bars + 2 * kConsumers + kStages);  // [kConsumers]: stage | record << 8 | half << 16
  uint8_t* recring = reinterpret_cast<uint8_t*>(bars) + kBarBytes;                           // [2][32] EdgeRec
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
    for (int w = 0; w < kConsumers; w++) { mbar_init(full0 + 8 * w, 1); mbar_init(taken0 + 8 * w, 1); }
    for (int s = 0; s < kStages; s++) mbar_init(empty0 + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  const long long BE = (long long)P.B * P.E;
  const long long lo = BE * blockIdx.x / gridDim.x, hi = BE * (blockIdx.x + 1) / gridDim.x;
  const int nt = (int)(hi - lo);
  const int per = (nt + kConsumers - 1) / kConsumers;  // edges per consumer warp (contiguous runs: consecutive edges share patches)

  if (warp >= kConsumers) {
    // ================================================================= producers
    // producer pi serves consumer warps [pi*WPP, (pi+1)*WPP) through stages [pi*SPP, (pi+1)*SPP); lanes 0 .. 4*WPP-1 work
    constexpr int WPP = kConsumers / kProducers, SPP = kStages / kProducers;
    static_assert(kConsumers % kProducers == 0 && kStages % kProducers == 0 && 4 * WPP <= 32, "producer split");
    const int pi = warp - kConsumers;
    const bool plane = lane < 4 * WPP;
    const int pw = pi * WPP + ((lane >> 2) % WPP), pj = lane & 3;  // this lane prepares AND issues edge 4*batch + pj of consumer warp pw
    const int cntw = plane ? max(0, min(per, nt - pw * per)) : 0;
    long long carry_pid = -1;                            // patch of warp pw's last edge of the previous batch
    int m_mine = 0;                                      // slots issued so far to consumer warp pw (same in its four lanes)
    int seq = 0;                                         // slots issued so far by the whole warp (uniform)
    const int nbatch = (per + 3) >> 2;
    for (int bt = 0; bt < nbatch; bt++) {
      // ---- every lane prepares one edge; the 32 edges' global loads overlap
      const int k = 4 * bt + pj;
      const bool valid = k < cntw;
      int kind = 3, x0 = 0, y0 = 0, slot = 0, has_patch = 0;
      long long pid = -1;
      const int recidx = (pi * 2 + (bt & 1)) * (4 * WPP) + lane;
      EdgeRec* myrec = reinterpret_cast<EdgeRec*>(recring + recidx * kRecBytes);
      if (valid) {
        const long long be = lo + (long long)pw * per + k;
        const int b = (int)(be / P.E), e = (int)(be % P.E);
        const int64_t ix = P.ii[e], jx = P.jj[e];
        const float2* c2 = reinterpret_cast<const float2*>(P.coords + be * (2 * kPP));  // 72-B records: 8-byte aligned
        float v[2 * kPP];
#pragma unroll
        for (int i = 0; i < kPP; i++) { const float2 t = c2[i]; v[2 * i] = t.x; v[2 * i + 1] = t.y; }
        int fx[kPP], fy[kPP];
        int mnx = kClamp * 2, mxx = -kClamp * 2, mny = kClamp * 2, mxy = -kClamp * 2;
#pragma unroll
        for (int p = 0; p < kPP; p++) {
          fx[p] = clamp_floor(v[p]); fy[p] = clamp_floor(v[kPP + p]);
          myrec->dx[p] = v[p] - floorf(v[p]); myrec->dy[p] = v[kPP + p] - floorf(v[kPP + p]);
          mnx = min(mnx, fx[p]); mxx = max(mxx, fx[p]); mny = min(mny, fy[p]); mxy = max(mxy, fy[p]);
        }
        x0 = mnx - kR; y0 = mny - kR;
        const int wdt = mxx - mnx + kD, hgt = mxy - mny + kD;
        const bool bad = ix < 0 || ix >= P.n1 || jx < 0 || jx >= P.n2;
        const bool outside = x0 >= P.W || y0 >= P.H || x0 + wdt <= 0 || y0 + hgt <= 0;
        if (bad) set_status(P.status, WV_STATUS_BAD_INDEX);
        int BX = kD;
        if (bad || outside) kind = 3;
        else if (wdt <= kBoxSmall && hgt <= kBoxSmall) { kind = 0; BX = kBoxSmall; }
        else if (wdt <= kBoxLarge && hgt <= kBoxLarge) { kind = 1; BX = kBoxLarge; }
        else kind = 2;
        slot = bad ? 0 : b * P.n2 + (int)jx;
        if (kind <= 2) pid = (long long)b * P.n1 + ix;
#pragma unroll
        for (int p = 0; p < kPP; p++) myrec->org[p] = (unsigned char)(kind <= 1 ? (fy[p] - mny) * BX + (fx[p] - mnx) : 0);
      }
      {
        const long long up = __shfl_up_sync(0xffffffffu, pid, 1);
        const long long prev = pj == 0 ? carry_pid : up;
        has_patch = (pid >= 0 && (pid != prev || kind == 2)) ? 1 : 0;
        carry_pid = __shfl_sync(0xffffffffu, pid, (lane & ~3) | 3);
      }
      if (valid) myrec->flags = kind | (has_patch << 2);
      const unsigned vmask = __ballot_sync(0xffffffffu, valid), lmask = __ballot_sync(0xffffffffu, valid && kind == 1);
      __syncwarp();

      // ---- the ring is fed in a fixed order (edge j, half, warp w): warps round-robin, each warp's slots in order.
      // The lane that prepared an edge issues it itself (its operands are in its own registers).
      for (int jp = 0; jp < 8; jp++) {
        const int j = jp >> 1, part = jp & 1;
        unsigned sub = ((part == 0 ? vmask : lmask) >> j) & 0x11111111u;
        while (sub) {
          const int bit = __ffs(sub) - 1;
          sub &= sub - 1;
          const int w = pi * WPP + (bit >> 2), src = bit + j;
          const int stage = pi * SPP + seq % SPP;
          if (lane == src) {
            mbar_wait(empty0 + 8 * stage, ((seq / SPP) & 1) ^ 1);
            if (m_mine > 0) mbar_wait(taken0 + 8 * w, (m_mine - 1) & 1);
            mailbox[w] = stage | (recidx << 8) | (part << 16);
            const uint32_t bar = full0 + 8 * w, dst = smem_u32(stages + (size_t)stage * kStageBytes);
            const bool patch_now = has_patch && part == 0;
            const int npix = kind == 0 ? kBoxSmall * kBoxSmall : kBoxLarge * kBoxLargeRows;
            const uint32_t bytes = (kind <= 1 ? 2u * npix * 128u : 0u) + (patch_now ? (uint32_t)kPatchBytes : 0u);
            if (bytes) mbar_arrive_expect_tx(bar, bytes); else mbar_arrive(bar);
            if (kind <= 1) {
              const CUtensorMap* map = kind == 0 ? &map_small : &map_half;
              const int yy = y0 + part * kBoxLargeRows;
              tma_load_4d(dst, map, bar, 0, x0, yy, slot);
              tma_load_4d(dst + kHalfBytes, map, bar, 64, x0, yy, slot);
            }
            if (patch_now) bulk_copy_g2s(dst + kPatchOff, P.gT + pid * (kC * kPP), kPatchBytes, bar);
          }
          if (w == pw) m_mine++;
          seq++;
          __syncwarp();
        }
      }
    }
    return;
  }

  // =================================================================== consumers (no global reads)
  float* rawD = reinterpret_cast<float*>(raw_area + (size_t)warp * kRawBytes);  // [9][kRawStride]: cols 0..143 D, 144..146 geometry
  const int g8 = lane >> 2, t4 = lane & 3;
  const uint32_t lds_off0 = (uint32_t)((lane & 7) * 128 + ((((lane >> 3) ^ (lane & 7)) & 7) << 4));  // ldmatrix row + swizzled chunk
  const uint32_t lds_off1 = lds_off0 ^ 64u;                                                            // chunks 4..7 of the 128-B row
  // A-fragment ldmatrix address inside the [9][128] patch: matrices (rows 0-7 | 8-15) x (k 0-7 | 8-15); rows > 8 alias row 8
  const uint32_t a_off = (uint32_t)(min(((lane >> 3) & 1) * 8 + (lane & 7), 8) * 256 + (lane >> 4) * 16);
  float* const d0 = rawD + g8 * kRawStride + 2 * t4;
  float* const d8 = rawD + 8 * kRawStride + 2 * t4;
  const bool has8 = g8 == 0;
  uint32_t A[8][4];
#pragma unroll
  for (int ks = 0; ks < 8; ks++) { A[ks][0] = 0u; A[ks][1] = 0u; A[ks][2] = 0u; A[ks][3] = 0u; }
  const int cnt = max(0, min(per, nt - warp * per));
  const uint32_t fullb = full0 + 8 * warp, takenb = taken0 + 8 * warp;

  int m = 0;
  for (int done = 0; done < cnt;) {
    mbar_wait(fullb, m & 1);
    const int mb = mailbox[warp];
    const int stage = mb & 0xff, part = (mb >> 16) & 1;
    const uint8_t* sb = stages + (size_t)stage * kStageBytes;
    const EdgeRec* rec = reinterpret_cast<const EdgeRec*>(recring + ((mb >> 8) & 0xff) * kRecBytes);
    const int flags = rec->flags, kind = flags & 3, has_patch = part == 0 ? (flags >> 2) & 1 : 0;
    const int nparts = kind == 1 ? 2 : 1, BX = kind == 0 ? kBoxSmall : (kind == 1 ? kBoxLarge : kD);
    const long long be = lo + (long long)warp * per + done;
    if (part == 0 && lane < kPP) {  // geometry of each patch pixel next to its D row
      float* gr = rawD + lane * kRawStride + kMaxPix;
      gr[0] = __int_as_float((int)rec->org[lane]); gr[1] = rec->dx[lane]; gr[2] = rec->dy[lane];
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(takenb);  // the record and the mailbox may be reused from here on
    m++;
    const uint32_t st = smem_u32(sb);

    if (has_patch) {
#pragma unroll
      for (int ks = 0; ks < 8; ks++) ldmatrix_x4(st + kPatchOff + a_off + ks * 32, A[ks][0], A[ks][1], A[ks][2], A[ks][3]);
    }

    if (kind <= 1) {
      // -------------------------------------------------------------- tensor-core contraction from the TMA tile
      const uint32_t a0 = st + lds_off0, a1 = st + lds_off1;
      if (kind == 0) {  // 100 px -> 13 pixel tiles (the last one half used)
        mma_group<7>(a0, a1, A, d0, d8, has8);  // 7 + 6 tiles: two dependent chains of 8 MMAs instead of four
        mma_group<6>(a0 + 7 * 1024, a1 + 7 * 1024, A, d0 + 56, d8 + 56, has8);
      } else {          // half of a 12x12 window: 72 px -> 9 pixel tiles, D columns 72*part ..
        const int c0 = part * (kBoxLarge * kBoxLargeRows);
        mma_group<5>(a0, a1, A, d0 + c0, d8 + c0, has8);
        mma_group<4>(a0 + 5 * 1024, a1 + 5 * 1024, A, d0 + c0 + 40, d8 + c0 + 40, has8);
      }
    } else if (kind == 2) {
      // -------------------------------------------------------------- direct gather (windows too far apart for a tile)
      // (rare; this path alone reads global memory from a consumer: coords and jj of its edge)
      const size_t plane = (size_t)P.H * P.W;
      const int eb = (int)(be / P.E), ee = (int)(be % P.E);
      const __half* f2 = P.fmap2 + ((size_t)eb * P.n2 + (size_t)P.jj[ee]) * kC * plane;
      const __half* pt = reinterpret_cast<const __half*>(sb + kPatchOff);  // [9][128]
      const float cv = lane < 2 * kPP ? P.coords[be * (2 * kPP) + lane] : 0.f;
      const int fl = clamp_floor(cv);
      for (int t = lane; t < kPP * kD * kD; t += 32) {
        const int b = t & 7, a = (t >> 3) & 7, p = t >> 6;
        const int i1 = __shfl_sync(0xffffffffu, fl, kPP + p) + a - kR, j1 = __shfl_sync(0xffffffffu, fl, p) + b - kR;
        float sacc = 0.f;
        if (i1 >= 0 && i1 < P.H && j1 >= 0 && j1 < P.W) {
          const __half* p2 = f2 + (size_t)i1 * P.W + j1;
          for (int c = 0; c < kC; c++) sacc = fmaf(__half2float(pt[p * kC + c]), __half2float(p2[(size_t)c * plane]), sacc);
        }
        rawD[p * kRawStride + a * kD + b] = sacc;
      }
    } else {
      for (int t = lane; t < kPP * kD * kD; t += 32) rawD[(t >> 6) * kRawStride + (t & 63)] = 0.f;
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty0 + 8 * stage);  // record, patch and tile of this stage are no longer read
    if (part != nparts - 1) continue;

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
    done++;
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// pass 2, tcgen05 form: corr_tc5_kernel.  Same tiles, same ring, but the contraction runs on the 5th-generation tensor cores:
//   D[128 window pixels x 16] = tile[128 px x 128 ch] . patch[16 x 128 ch]^T        (100 or 72 pixel rows and 9 columns are real)
// with BOTH operands read straight from the 128-B-swizzled shared memory the TMA wrote (K-major canonical layout: the tile IS the
// A operand, M = window pixels; the pre-swizzled patch is the B operand, N = 16), the accumulator in TMEM, 8 tcgen05.mma
// (M128 N16 K16) per tile issued by ONE thread.  mma.sync needs 104 warp-wide m16n8k16 per tile at ~16 cycles each on a
// sub-core's legacy path (66 us per level for the tensor pipe alone, 7/16 of it on padding rows); here the MMA cost vanishes
// and, because tcgen05.commit releases a stage the moment its MMAs retire, stages recycle without waiting for the blend.
//
// Two independent pipelines per CTA (each half of the CTA's edges), 6 warps each:
//   producer warp  16 lanes prepare an edge each (window box, per-pixel origins + fractions -> a record), then issue in order:
//                  wait empty[stage] -> expect_tx on full[stage] -> 2 TMA tile loads + 2 bulk copies of the patch halves
//   MMA warp       one lane: wait full[stage], wait accempty[acc], 8 x tcgen05.mma, commit -> empty[stage], commit -> accfull[acc]
//   4 consumer warps (one TMEM lane quarter each): wait accfull[acc], tcgen05.ld 32 pixel rows x 16 columns, scatter the 9 real
//                  columns into the edge's D scratch, release the accumulator; after the edge's last part a 128-thread barrier,
//                  then the 441 outputs of the edge, 4-tap bilinear each, written as coalesced fp16.
// Every mbarrier has one waiter (or one group waiting for the same phase) and advances strictly in slot order.
// ---------------------------------------------------------------------------------------------------
namespace {

constexpr int kT5Pipes = 3, kT5Stages = 2, kT5Acc = 4, kT5Rec = 32;
constexpr int kT5RawStride = 161;                             // chosen by a bank simulation of the blend's reads on tracker geometry (2 wavefronts per load)
constexpr int kT5RawBytes = kPP * kT5RawStride * 4;
constexpr int kT5Warps = 6;                                   // per pipeline: producer, MMA, 4 consumers
constexpr int kT5Threads = 32 * kT5Pipes * kT5Warps;          // 384
constexpr int kT5StageBytes = 30 * 1024;                      // two 13 KB tile halves + the patch as B operand [2][16][128 B]
constexpr int kT5PatchOff = 26 * 1024, kT5PatchHalf = 2048, kT5PatchCopy = kPP * 128;  // 9 real rows per half
constexpr int kT5RecBytes = 104;  // sizeof(T5Rec), 8-byte aligned slots
constexpr int kT5PipeBytes = kT5Stages * kT5StageBytes + 2 * kT5RawBytes + kT5Rec * kT5RecBytes;
constexpr int kT5Bars = 2 * kT5Stages + 2 * kT5Acc;           // per pipeline
constexpr int kT5Smem = 1024 + kT5Pipes * kT5PipeBytes + kT5Pipes * kT5Bars * 8 + 64;
static_assert(kT5Smem <= 227 * 1024 && kT5Pipes * kT5Acc * 16 <= 256, "tcgen05 kernel shared memory / TMEM budget");
static_assert((kT5Stages * kT5StageBytes) % 1024 == 0 && kT5PipeBytes % 8 == 0 && kMaxPix + 3 <= kT5RawStride, "alignment");

struct T5Rec {  // one per SLOT (an edge with a 12x12 window owns two)
  int kind;                  // 0 small tile, 1 half of a large tile, 2 direct gather, 3 nothing to read, 4 end of stream
  short part, nparts;
  int level;                 // pyramid level of the slot (two-level form)
  int be;                    // b * E + e
  unsigned char org[12];
  float dx[kPP], dy[kPP];
};

__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {  // K-major, 128-B swizzle, 8-row groups 1024 B apart
  return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

static_assert(sizeof(T5Rec) <= kT5RecBytes, "record slot");

// NL = 1: one map ring, out [B][E][7][7][3][3].  NL = 2 (the tracker's DPVO.corr, dpvo/dpvo.py:325-333): every edge goes through
// the pipeline twice in a row -- level 0 against fmap2 at coords * cmul0, then level 1 against fmap2_l1 at coords * cmul1 -- and
// the consumer that owns an output keeps its level-0 value in a register until the level-1 value exists, then stores the pair:
// out [B][E][7][7][3][3][2] == torch.stack([corr1, corr2], -1), written once, 4 bytes per thread.
template <int NL>
__global__ void __launch_bounds__(kT5Threads, 1)
corr_tc5_kernel(const __grid_constant__ CorrMaps maps, const CorrParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint32_t tmem_base_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int pl = warp / kT5Warps, role = warp % kT5Warps;
  // stages of both pipelines first (1024-B aligned), then per pipeline: D scratch x2, records; barriers last
  uint8_t* stages = smem + (size_t)pl * kT5Stages * kT5StageBytes;
  uint8_t* aux = smem + (size_t)kT5Pipes * kT5Stages * kT5StageBytes + (size_t)pl * (2 * kT5RawBytes + kT5Rec * kT5RecBytes);
  float* raw0 = reinterpret_cast<float*>(aux);
  uint8_t* recs = aux + 2 * kT5RawBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kT5Pipes * kT5Stages * kT5StageBytes +
                                               (size_t)kT5Pipes * (2 * kT5RawBytes + kT5Rec * kT5RecBytes)) + pl * kT5Bars;
  const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * kT5Stages, accfull0 = empty0 + 8 * kT5Stages,
                 accempty0 = accfull0 + 8 * kT5Acc;

  if (role == 0 && lane == 0) {
    for (int i = 0; i < kT5Stages; i++) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
    for (int i = 0; i < kT5Acc; i++) { mbar_init(accfull0 + 8 * i, 1); mbar_init(accempty0 + 8 * i, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s + (uint32_t)(pl * kT5Acc * 16);  // this pipeline's accumulators: 16 columns each

  const long long BE = (long long)P.B * P.E;
  const long long lo = BE * blockIdx.x / gridDim.x, hi = BE * (blockIdx.x + 1) / gridDim.x;
  const long long plo = lo + (hi - lo) * pl / kT5Pipes, phi = lo + (hi - lo) * (pl + 1) / kT5Pipes;
  const int cnt = (int)(phi - plo);

  if (role == 0) {
    // ================================================================= producer
    int sbase = 0;  // slots issued so far (warp-uniform)
    const int nitems = NL * cnt;                      // item = (edge, level), an edge's levels back to back
    for (int kbase = 0;;) {
      const int k = kbase + lane;
      const bool last = kbase >= nitems;              // the batch after the last one carries the end-of-stream slot in lane 0
      const bool valid = !last && lane < 16 && k < nitems;
      const int level = NL == 1 ? 0 : (k & 1);
      const int LH = level ? P.H1 : P.H, LW = level ? P.W1 : P.W;
      int kind = 3, x0 = 0, y0 = 0, slot = 0, nsl = 0, bxs = kD;
      long long pid = -1, be = 0;
      unsigned char org[12];
      float dxv[kPP], dyv[kPP];
      if (valid) {
        be = plo + (NL == 1 ? k : (k >> 1));
        const int b = (int)(be / P.E), e = (int)(be % P.E);
        const int64_t ix = P.ii[e], jx = P.jj[e];
        const float2* c2 = reinterpret_cast<const float2*>(P.coords + be * (2 * kPP));
        float v[2 * kPP];
#pragma unroll
        for (int i = 0; i < kPP; i++) { const float2 t = c2[i]; v[2 * i] = t.x; v[2 * i + 1] = t.y; }
        if (NL == 2) {  // dpvo/dpvo.py:331-332: coords / 1, coords / 4 (multipliers are exact powers of two)
          const float cm = level ? P.cmul1 : P.cmul0;
#pragma unroll
          for (int i = 0; i < 2 * kPP; i++) v[i] *= cm;
        }
        int fx[kPP], fy[kPP];
        int mnx = kClamp * 2, mxx = -kClamp * 2, mny = kClamp * 2, mxy = -kClamp * 2;
#pragma unroll
        for (int p = 0; p < kPP; p++) {
          fx[p] = clamp_floor(v[p]); fy[p] = clamp_floor(v[kPP + p]);
          dxv[p] = v[p] - floorf(v[p]); dyv[p] = v[kPP + p] - floorf(v[kPP + p]);
          mnx = min(mnx, fx[p]); mxx = max(mxx, fx[p]); mny = min(mny, fy[p]); mxy = max(mxy, fy[p]);
        }
        x0 = mnx - kR; y0 = mny - kR;
        const int wdt = mxx - mnx + kD, hgt = mxy - mny + kD;
        const bool bad = ix < 0 || ix >= P.n1 || jx < 0 || jx >= P.n2;
        const bool outside = x0 >= LW || y0 >= LH || x0 + wdt <= 0 || y0 + hgt <= 0;
        if (bad) set_status(P.status, WV_STATUS_BAD_INDEX);
        int BX = kD;
        if (bad || outside) kind = 3;
        else if (wdt <= kBoxTiny && hgt <= kBoxTiny) { kind = 0; BX = kBoxTiny; }
        else if (wdt <= kBoxSmall && hgt <= kBoxSmall) { kind = 0; BX = kBoxSmall; }
        else if (wdt <= kBoxLarge && hgt <= kBoxLarge) { kind = 1; BX = kBoxLarge; }
        else kind = 2;
        bxs = BX;
        slot = bad ? 0 : b * P.n2 + (int)jx;
        if (kind <= 2) pid = (long long)b * P.n1 + ix;
#pragma unroll
        for (int p = 0; p < kPP; p++) org[p] = (unsigned char)(kind <= 1 ? (fy[p] - mny) * BX + (fx[p] - mnx) : 0);
        nsl = kind == 1 ? 2 : 1;
      }
      if (last && lane == 0) { kind = 4; nsl = 1; }
      // slot numbers: exclusive prefix sum of nsl over the lanes
      int incl = nsl;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
      // Records of the last 6 slots issued may still be in use (2 stages + 4 accumulators in flight): a batch writes at most
      // kT5Rec - 6 records; the lanes beyond that (only when most of the 16 windows need two slots) go to the next batch.
      const unsigned cutmask = __ballot_sync(0xffffffffu, incl > kT5Rec - 6);
      const int kept = cutmask ? __ffs(cutmask) - 1 : 32;   // >= 13
      if (lane >= kept) nsl = 0;
      const int myslot = sbase + incl - nsl;
      const int total = __shfl_sync(0xffffffffu, incl, kept - 1);
      for (int part = 0; part < nsl; part++) {
        T5Rec* r = reinterpret_cast<T5Rec*>(recs + ((myslot + part) % kT5Rec) * kT5RecBytes);
        r->kind = kind; r->part = (short)part; r->nparts = (short)nsl; r->level = level | (bxs << 8); r->be = (int)be;
#pragma unroll
        for (int p = 0; p < kPP; p++) { r->org[p] = org[p]; r->dx[p] = dxv[p]; r->dy[p] = dyv[p]; }
      }
      __syncwarp();
      // ---- issue in slot order; the lane that prepared an edge issues it (its operands are in its registers)
      for (int L = 0; L < 16; L++) {
        const int nL = __shfl_sync(0xffffffffu, nsl, L);
        if (lane == L) {
          for (int part = 0; part < nL; part++) {
            const int sidx = myslot + part, stage = sidx % kT5Stages;
            mbar_wait(empty0 + 8 * stage, ((sidx / kT5Stages) & 1) ^ 1);
            const uint32_t bar = full0 + 8 * stage, dst = smem_u32(stages + (size_t)stage * kT5StageBytes);
            const int npix = kind == 0 ? bxs * bxs : kBoxLarge * kBoxLargeRows;
            const uint32_t bytes = (kind <= 1 ? 2u * npix * 128u : 0u) + (kind <= 2 ? 2u * kT5PatchCopy : 0u);
            if (bytes) mbar_arrive_expect_tx(bar, bytes); else mbar_arrive(bar);
            if (kind <= 1) {
              const CUtensorMap* map = &maps.m[3 * level + (kind == 1 ? 1 : (bxs == kBoxTiny ? 2 : 0))];
              const int yy = y0 + part * kBoxLargeRows;
              tma_load_4d(dst, map, bar, 0, x0, yy, slot);
              tma_load_4d(dst + kHalfBytes, map, bar, 64, x0, yy, slot);
            }
            if (kind <= 2) {
              const __half* src = P.gT + pid * (kC * kPP);
              bulk_copy_g2s(dst + kT5PatchOff, src, kT5PatchCopy, bar);
              bulk_copy_g2s(dst + kT5PatchOff + kT5PatchHalf, src + kT5PatchCopy / 2, kT5PatchCopy, bar);
            }
          }
        }
        __syncwarp();
      }
      sbase += total;
      if (last) break;
      kbase += kept < 16 ? kept : 16;
    }
  } else if (role == 1) {
    // ================================================================= MMA issue (one lane)
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | ((16u >> 3) << 17) | ((128u >> 4) << 24);  // D fp32; A, B fp16 K-major; N = 16; M = 128
      for (int sidx = 0;; sidx++) {
        const int stage = sidx % kT5Stages, acc = sidx % kT5Acc;
        mbar_wait(full0 + 8 * stage, (sidx / kT5Stages) & 1);
        const int kind = reinterpret_cast<const T5Rec*>(recs + (sidx % kT5Rec) * kT5RecBytes)->kind;
        mbar_wait(accempty0 + 8 * acc, ((sidx / kT5Acc) & 1) ^ 1);
        if (kind <= 1) {
          tc_fence_after();
          const uint32_t sa = smem_u32(stages + (size_t)stage * kT5StageBytes);
          const uint64_t da0 = umma_desc(sa), db0 = umma_desc(sa + kT5PatchOff);  // advancing K or the half only moves the start address
#pragma unroll
          for (int kq = 0; kq < 8; kq++) {
            const uint32_t h = kq >> 2, kk = kq & 3;
            umma_f16(tmem + acc * 16, da0 + ((h * kHalfBytes + kk * 32) >> 4), db0 + ((h * kT5PatchHalf + kk * 32) >> 4), idesc,
                     kq > 0 ? 1u : 0u);
          }
          umma_commit(empty0 + 8 * stage);   // the stage is free once these MMAs have read it
          umma_commit(accfull0 + 8 * acc);   // ... and the accumulator is complete
        } else {
          if (kind != 2) mbar_arrive(empty0 + 8 * stage);  // (a gather edge keeps its stage: the consumers read the patch)
          mbar_arrive(accfull0 + 8 * acc);
          if (kind == 4) break;
        }
      }
    }
  } else {
    // ================================================================= consumers: one warpgroup per pipeline
    const int q = warp & 3;                       // the TMEM lane quarter this warp may read
    const int wt = (role - 2) * 32 + lane;        // thread index in the warpgroup
    // outputs of an edge: out[x-tap b][y-tap a][p].  Thread wt < 126 owns patch pixel p, y-tap a and the x-taps 4 bh .. 4 bh + 3
    // (bh = 1: three of them): it reads 2 rows x 5 adjacent columns of p's window once (10 loads for 4 outputs instead of 16), and
    // for a fixed x-tap the 63 threads (a, p) write 63 consecutive outputs.  The kernel is bound by shared-memory bandwidth (TMA
    // writes + UMMA operand reads + this epilogue), so loads saved here are time saved.
    const int o_bh = wt / 63, o_r = wt - 63 * o_bh, o_a = o_r / kPP, o_p = o_r - kPP * o_a;
    int eidx = 0;
    __half keep[4];                               // two-level form: this thread's level-0 outputs of the edge in flight
#pragma unroll
    for (int j = 0; j < 4; j++) keep[j] = __float2half_rn(0.f);
    for (int sidx = 0;; sidx++) {
      const int acc = sidx % kT5Acc, stage = sidx % kT5Stages;
      mbar_wait(accfull0 + 8 * acc, (sidx / kT5Acc) & 1);
      const T5Rec* rec = reinterpret_cast<const T5Rec*>(recs + (sidx % kT5Rec) * kT5RecBytes);
      const int kind = rec->kind;
      if (kind == 4) break;
      const int part = rec->part, nparts = rec->nparts, lvbx = rec->level, level = NL == 1 ? 0 : (lvbx & 1), bxs = lvbx >> 8;
      const long long be = rec->be;
      float* rawD = raw0 + (eidx & 1) * (kT5RawBytes / 4);
      if (part == 0 && wt < kPP) {
        float* gr = rawD + wt * kT5RawStride + kMaxPix;
        gr[0] = __int_as_float((int)rec->org[wt]); gr[1] = rec->dx[wt]; gr[2] = rec->dy[wt];
      }
      if (kind <= 1) {
        tc_fence_after();
        uint32_t v[16];
        const uint32_t taddr = tmem + acc * 16 + ((uint32_t)(32 * q) << 16);
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
              "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
            : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const int row = 32 * q + lane, npix = kind == 0 ? bxs * bxs : kBoxLarge * kBoxLargeRows;
        if (row < npix) {
          float* d = rawD + part * (kBoxLarge * kBoxLargeRows) + row;
#pragma unroll
          for (int p = 0; p < kPP; p++) d[p * kT5RawStride] = __uint_as_float(v[p]);
        }
        tc_fence_before();
      } else if (kind == 2) {
        // direct gather (windows too far apart for a tile): the only path that reads global memory from a consumer
        const int LH = level ? P.H1 : P.H, LW = level ? P.W1 : P.W;
        const float cm = NL == 1 ? 1.f : (level ? P.cmul1 : P.cmul0);
        const size_t plane = (size_t)LH * LW;
        const int eb = (int)(be / P.E), ee = (int)(be % P.E);
        const __half* f2 = (level ? P.fmap2_l1 : P.fmap2) + ((size_t)eb * P.n2 + (size_t)P.jj[ee]) * kC * plane;
        const uint8_t* pb = stages + (size_t)stage * kT5StageBytes + kT5PatchOff;
        for (int t = wt; t < kPP * kD * kD; t += 128) {
          const int bb = t & 7, a = (t >> 3) & 7, p = t >> 6;
          const int i1 = clamp_floor(P.coords[be * (2 * kPP) + kPP + p] * cm) + a - kR,
                    j1 = clamp_floor(P.coords[be * (2 * kPP) + p] * cm) + bb - kR;
          float sacc = 0.f;
          if (i1 >= 0 && i1 < LH && j1 >= 0 && j1 < LW) {
            const __half* p2 = f2 + (size_t)i1 * LW + j1;
            for (int c = 0; c < kC; c++) {
              const __half pv = *reinterpret_cast<const __half*>(pb + (c >> 6) * kT5PatchHalf + p * 128 + (((((c & 63) >> 3) ^ (p & 7)) & 7) << 4) + (c & 7) * 2);
              sacc = fmaf(__half2float(pv), __half2float(p2[(size_t)c * plane]), sacc);
            }
          }
          rawD[p * kT5RawStride + a * kD + bb] = sacc;
        }
      } else {
        for (int t = wt; t < kPP * kD * kD; t += 128) rawD[(t >> 6) * kT5RawStride + (t & 63)] = 0.f;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(accempty0 + 8 * acc);
      if (part != nparts - 1) continue;

      asm volatile("bar.sync %0, 128;" ::"r"(1 + pl) : "memory");  // the edge's D scratch is complete
      if (kind == 2 && wt == 0) mbar_arrive(empty0 + 8 * stage);
      // ---- the 441 outputs of the edge: out[x-tap b][y-tap a][p], 4-tap bilinear from pixel p's own 8x8 window
      const int BX = kind <= 1 ? bxs : kD;
      __half* o = P.out + be * (NL * kT * kT * kPP);
      if (wt < 2 * 63) {
        const float* row = rawD + o_p * kT5RawStride;
        const float dx = row[kMaxPix + 1], dy = row[kMaxPix + 2];
        const float* rr = row + __float_as_int(row[kMaxPix]) + o_a * BX + 4 * o_bh;
        float t0[5], t1[5];
#pragma unroll
        for (int c = 0; c < 4; c++) { t0[c] = rr[c]; t1[c] = rr[BX + c]; }
        if (o_bh == 0) { t0[4] = rr[4]; t1[4] = rr[BX + 4]; } else { t0[4] = 0.f; t1[4] = 0.f; }   // column 8 is outside the window
#pragma unroll
        for (int j = 0; j < 4; j++) {
          if (j < 3 || o_bh == 0) {
            const float h0 = fmaf(dx, t0[j + 1] - t0[j], t0[j]), h1 = fmaf(dx, t1[j + 1] - t1[j], t1[j]);
            const __half val = __float2half_rn(fmaf(dy, h1 - h0, h0));
            const int idx = (4 * o_bh + j) * 63 + o_r;
            if (NL == 1) o[idx] = val;
            else if (level == 0) keep[j] = val;
            else reinterpret_cast<__half2*>(o)[idx] = __halves2half2(keep[j], val);  // (level 0, level 1) pair
          }
        }
      }
      eidx++;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem_base_s) : "memory");
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cudaFuncAttributeMaxDynamicSharedMemorySize is per device: remember per (kernel, device) whether it has been raised, so that one
// process driving several GPUs (clip batches in threads) opts in on each of them.  Returns the previous state and marks it set.
bool smem_opted_in(int kernel) {
  static bool done[3][64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return false;   // unknown device: set the attribute every call
  const bool was = done[kernel][dev];
  done[kernel][dev] = true;
  return was;
}

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

bool make_map(CUtensorMap* m, void* nhwc, int W, int H, long long slots, int box_x, int box_y) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  const cuuint64_t dims[4] = {(cuuint64_t)kC, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)slots};
  const cuuint64_t strides[3] = {(cuuint64_t)kC * 2, (cuuint64_t)kC * 2 * W, (cuuint64_t)kC * 2 * W * H};
  const cuuint32_t boxd[4] = {64, (cuuint32_t)box_x, (cuuint32_t)box_y, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, nhwc, dims, strides, boxd, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int tiled_mode() {  // WVB200_CORR_TILED: unset = heuristic, 0 = never, 1 = whenever the shape allows
  const char* s = getenv("WVB200_CORR_TILED");
  if (!s || !s[0]) return -1;
  return s[0] == '0' ? 0 : 1;
}

inline size_t al1k(size_t x) { return (x + 1023) & ~(size_t)1023; }

bool tc5_mode() {  // WVB200_CORR_TC5=0: the mma.sync kernel (corr_tiled_kernel) instead of the tcgen05 one
  const char* s = getenv("WVB200_CORR_TC5");
  return !(s && s[0] == '0');
}

}  // namespace

bool corr_tiled_supported(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int R) {
  const int mode = tiled_mode();
  if (mode == 0) return false;
  if (dtype != WV_F16 || C != kC || PH != kP || PW != kP || R != kR) return false;
  if (B <= 0 || E <= 0 || n1 <= 0 || (long long)B * n2 > 0x7fffffffLL || H2 > (1 << 20) || W2 > (1 << 20)) return false;
  if (!encode_tiled_fn()) return false;
  if (mode == 1) return true;
  // restaging costs one read + one write of the ring and of the patches: worth it once the edges' tile traffic dominates
  const double ring = ((double)B * n2 * H2 * W2 + (double)B * n1 * kPP) * kC * 2.0;
  const double tiles = (double)B * E * 2.0 * kBoxSmall * kBoxSmall * 128.0;
  return (long long)B * E >= 1024 && tiles >= 0.5 * ring;
}

size_t corr_tiled_workspace_bytes(int, int B, int, int, int, int, int n1, int n2, int H2, int W2, int) {
  return 1024 + al1k((size_t)B * n2 * H2 * W2 * kC * 2) + al1k((size_t)B * n1 * kPP * kC * 2);
}

int corr_forward_tiled(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii, const int64_t* jj, void* out,
                       int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int R, void* workspace,
                       size_t workspace_bytes, int* status, cudaStream_t st) {
  if (!corr_tiled_supported(dtype, B, E, C, PH, PW, n1, n2, H2, W2, R)) return WV_EUNSUPPORTED;
  if (workspace_bytes < corr_tiled_workspace_bytes(dtype, B, E, C, PH, PW, n1, n2, H2, W2, R)) return WV_EWORKSPACE;
  if ((reinterpret_cast<uintptr_t>(fmap1) & 1) || (reinterpret_cast<uintptr_t>(fmap2) & 1) ||
      (reinterpret_cast<uintptr_t>(coords) & 7))
    return WV_EUNSUPPORTED;
  const long long slots = (long long)B * n2, n_patches = (long long)B * n1;
  const int HW = H2 * W2;
  __half* nhwc = reinterpret_cast<__half*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~(uintptr_t)1023);
  __half* gT = reinterpret_cast<__half*>(reinterpret_cast<char*>(nhwc) + al1k((size_t)slots * HW * kC * 2));
  CUtensorMap ms, mh;
  if (!make_map(&ms, nhwc, W2, H2, slots, kBoxSmall, kBoxSmall) || !make_map(&mh, nhwc, W2, H2, slots, kBoxLarge, kBoxLargeRows))
    return WV_EUNSUPPORTED;

  const bool use_tc5 = tc5_mode();
  const int tiles_per_slot = (HW + kTrPix - 1) / kTrPix;
  const long long n_tiles = slots * tiles_per_slot;
  const long long cap = (long long)sm_count() * 8;
  const int tile_blocks = (int)(n_tiles < cap ? n_tiles : cap);
  const long long pb = (n_patches + 7) / 8;
  const int patch_blocks = (int)(pb < sm_count() ? pb : sm_count());
  const int vec_ok = ((HW & 7) == 0 && (reinterpret_cast<uintptr_t>(fmap2) & 15) == 0) ? 1 : 0;
  restage_kernel<<<tile_blocks + patch_blocks, 256, 0, st>>>((const __half*)fmap2, nhwc, HW, tiles_per_slot, n_tiles, vec_ok,
                                                            (const __half*)fmap1, gT, n_patches, tile_blocks, use_tc5 ? 1 : 0);
  WV_LAUNCH_CHECK();

  CorrParams P{gT, (const __half*)fmap2, coords, ii, jj, (__half*)out, B, E, n1, n2, H2, W2, status, nullptr, 0, 0, 1.f, 1.f};
  const long long BE = (long long)B * E;
  if (use_tc5) {
    if (!smem_opted_in(0)) WV_CUDA_TRY(cudaFuncSetAttribute(corr_tc5_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kT5Smem));
    CorrMaps maps;
    if (!make_map(&maps.m[2], nhwc, W2, H2, slots, kBoxTiny, kBoxTiny)) return WV_EUNSUPPORTED;
    maps.m[0] = ms; maps.m[1] = mh; maps.m[3] = ms; maps.m[4] = mh; maps.m[5] = maps.m[2];
    const long long want5 = (BE + 7) / 8;
    corr_tc5_kernel<1><<<(int)(want5 < sm_count() ? want5 : sm_count()), kT5Threads, kT5Smem, st>>>(maps, P);
    WV_LAUNCH_CHECK();
    return WV_OK;
  }
  if (!smem_opted_in(1)) WV_CUDA_TRY(cudaFuncSetAttribute(corr_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  const long long want = (BE + kConsumers - 1) / kConsumers;
  const int grid = (int)(want < sm_count() ? want : sm_count());
  corr_tiled_kernel<<<grid, kThreads, kSmemBytes, st>>>(ms, mh, P);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

// ---------------------------------------------------------------------------------------------------
// two-level (stacked) form: DPVO.corr (dpvo/dpvo.py:325-333) in one pass-2 launch.
//   out [B][E][7][7][3][3][2] fp16 == torch.stack([corr(gmap, pyramid[0], coords * cmul0), corr(gmap, pyramid[1], coords * cmul1)], -1)
// The reference runs 2 x (corr kernel + 9 ATen blend launches) and then torch.stack, which reads both 40 MB volumes and writes 80 MB.
// ---------------------------------------------------------------------------------------------------
bool corr_stacked_supported(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H0, int W0, int H1, int W1, int R) {
  if (dtype != WV_F16 || C != kC || PH != kP || PW != kP || R != kR) return false;
  if (B <= 0 || E <= 0 || n1 <= 0 || n2 <= 0 || (long long)B * n2 > 0x7fffffffLL || (long long)B * E > 0x7fffffffLL / 2) return false;
  if (H0 <= 0 || W0 <= 0 || H1 <= 0 || W1 <= 0 || H0 > (1 << 20) || W0 > (1 << 20) || H1 > (1 << 20) || W1 > (1 << 20)) return false;
  return encode_tiled_fn() != nullptr;
}

size_t corr_stacked_workspace_bytes(int B, int n1, int n2, int H0, int W0, int H1, int W1) {
  return 1024 + al1k((size_t)B * n2 * H0 * W0 * kC * 2) + al1k((size_t)B * n2 * H1 * W1 * kC * 2) + al1k((size_t)B * n1 * kPP * kC * 2);
}

int corr_forward_stacked(const void* fmap1, const void* fmap2_l0, const void* fmap2_l1, const float* coords, float cmul0, float cmul1,
                         const int64_t* ii, const int64_t* jj, void* out, int dtype, int B, int E, int C, int PH, int PW, int n1,
                         int n2, int H0, int W0, int H1, int W1, int R, void* workspace, size_t workspace_bytes, int* status,
                         cudaStream_t st) {
  if (!corr_stacked_supported(dtype, B, E, C, PH, PW, n1, n2, H0, W0, H1, W1, R)) return WV_EUNSUPPORTED;
  if (workspace_bytes < corr_stacked_workspace_bytes(B, n1, n2, H0, W0, H1, W1)) return WV_EWORKSPACE;
  if ((reinterpret_cast<uintptr_t>(fmap1) & 1) || (reinterpret_cast<uintptr_t>(fmap2_l0) & 1) ||
      (reinterpret_cast<uintptr_t>(fmap2_l1) & 1) || (reinterpret_cast<uintptr_t>(coords) & 7) || (reinterpret_cast<uintptr_t>(out) & 3))
    return WV_EUNSUPPORTED;
  const long long slots = (long long)B * n2, n_patches = (long long)B * n1;
  __half* nhwc0 = reinterpret_cast<__half*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~(uintptr_t)1023);
  __half* nhwc1 = reinterpret_cast<__half*>(reinterpret_cast<char*>(nhwc0) + al1k((size_t)slots * H0 * W0 * kC * 2));
  __half* gT = reinterpret_cast<__half*>(reinterpret_cast<char*>(nhwc1) + al1k((size_t)slots * H1 * W1 * kC * 2));
  CorrMaps maps;
  if (!make_map(&maps.m[0], nhwc0, W0, H0, slots, kBoxSmall, kBoxSmall) || !make_map(&maps.m[1], nhwc0, W0, H0, slots, kBoxLarge, kBoxLargeRows) ||
      !make_map(&maps.m[2], nhwc0, W0, H0, slots, kBoxTiny, kBoxTiny) ||
      !make_map(&maps.m[3], nhwc1, W1, H1, slots, kBoxSmall, kBoxSmall) || !make_map(&maps.m[4], nhwc1, W1, H1, slots, kBoxLarge, kBoxLargeRows) ||
      !make_map(&maps.m[5], nhwc1, W1, H1, slots, kBoxTiny, kBoxTiny))
    return WV_EUNSUPPORTED;

  // pass 1: level 0 together with the patches, then level 1
  for (int level = 0; level < 2; level++) {
    const void* src = level ? fmap2_l1 : fmap2_l0;
    const int HW = level ? H1 * W1 : H0 * W0;
    const int tiles_per_slot = (HW + kTrPix - 1) / kTrPix;
    const long long n_tiles = slots * tiles_per_slot;
    const long long cap = (long long)sm_count() * 8;
    const int tile_blocks = (int)(n_tiles < cap ? n_tiles : cap);
    const long long np = level ? 0 : n_patches;
    const long long pb = (np + 7) / 8;
    const int patch_blocks = (int)(pb < sm_count() ? pb : sm_count());
    const int vec_ok = ((HW & 7) == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0) ? 1 : 0;
    restage_kernel<<<tile_blocks + patch_blocks, 256, 0, st>>>((const __half*)src, level ? nhwc1 : nhwc0, HW, tiles_per_slot, n_tiles,
                                                              vec_ok, (const __half*)fmap1, gT, np, tile_blocks, 1);
    WV_LAUNCH_CHECK();
  }

  CorrParams P{gT, (const __half*)fmap2_l0, coords, ii, jj, (__half*)out, B, E, n1, n2, H0, W0, status,
               (const __half*)fmap2_l1, H1, W1, cmul0, cmul1};
  if (!smem_opted_in(2)) WV_CUDA_TRY(cudaFuncSetAttribute(corr_tc5_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kT5Smem));
  const long long want = ((long long)B * E + 7) / 8;
  corr_tc5_kernel<2><<<(int)(want < sm_count() ? want : sm_count()), kT5Threads, kT5Smem, st>>>(maps, P);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

}  // namespace wv
