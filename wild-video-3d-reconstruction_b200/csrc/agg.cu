// SoftAgg of the update operator (dpvo/blocks.py:31-48, called twice per update() from dpvo/net.py:88-89):
//
//     _, jx = torch.unique(ix, return_inverse=True)
//     w = torch_scatter.scatter_softmax(g(x), jx, dim=1)
//     y = torch_scatter.scatter_sum(f(x) * w, jx, dim=1)
//
// The reference runs a device sort + host sync (unique), then scatter_max, gather, sub, exp, scatter_sum, gather, div, mul and a
// second scatter_sum: ten passes over [E][384] tensors and two atomics-based scatters (torch-scatter 2.1.2, composite/softmax.py).
// Here:
//   group plan   the inverse index without a sort.  Key ranges up to 2^22 (the tracker's kk and ii*12345+jj) go through the
//                counting-sort edge plan shared with the BA (ba_plan.cu, one cooperative launch) with the key itself as the
//                secondary sort key: dense rank of a key among the present keys == torch.unique's sorted order == jx, segments ==
//                member lists in ascending edge order.  Wider ranges: presence bitmap -> popcount prefix -> rank, then counting
//                fill + a per-group rank sort (the kernels right below).  Either way the reduction order is fixed: deterministic.
//   soft_agg     ONE pass over f(x) and g(x): a CTA per group, a thread owns 16 bytes of channels and a share of the group's members
//                and walks them once in chunks of four (online softmax: running max, rescaled sums), fp32 arithmetic, one rounding
//                at the end.
//                HBM-bound: algorithmic bytes = 2 * E * D * sizeof(T) read + G * D * sizeof(T) written.
#include <cstdlib>

#include "ba_plan.cuh"
#include "wv_common.cuh"

namespace wv {

namespace {

constexpr int kScanThreads = 1024, kScanPer = 4, kScanTile = kScanThreads * kScanPer;

struct PopcOp {
  const uint32_t* w;
  __device__ __forceinline__ int operator()(long long i) const { return __popc(w[i]); }
};
struct CountOp {
  const int* c;
  long long n;  // entries beyond n read as 0 (the scan runs one past the end to produce the total)
  __device__ __forceinline__ int operator()(long long i) const { return i < n ? c[i] : 0; }
};

__device__ __forceinline__ int block_exclusive_scan(int v, int* total) {
  __shared__ int warp_sums[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int s = warp_sums[lane], si = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, si, o); if (lane >= o) si += t; }
    warp_sums[lane] = si - s;
    if (lane == 31 && total) *total = si;
  }
  __syncthreads();
  const int r = warp_sums[warp] + incl - v;
  __syncthreads();
  return r;
}

// phase A: per-tile sums
template <typename Op>
__global__ void __launch_bounds__(kScanThreads) scan_tile_sums(Op op, long long n, int* tile_sums) {
  const long long base = (long long)blockIdx.x * kScanTile + (long long)threadIdx.x * kScanPer;
  int v = 0;
#pragma unroll
  for (int i = 0; i < kScanPer; i++) if (base + i < n) v += op(base + i);
  __shared__ int tot;
  block_exclusive_scan(v, &tot);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

// phase B: one block turns the tile sums into exclusive tile offsets (any number of tiles: chunks of 1024 with a carry)
__global__ void __launch_bounds__(kScanThreads) scan_tile_offsets(int* tile_sums, int n_tiles, int* grand_total) {
  __shared__ int tot;
  int carry = 0;
  for (int c0 = 0; c0 < n_tiles; c0 += kScanThreads) {
    const int i = c0 + threadIdx.x;
    const int v = i < n_tiles ? tile_sums[i] : 0;
    const int ex = block_exclusive_scan(v, &tot);
    if (i < n_tiles) tile_sums[i] = carry + ex;
    carry += tot;
    __syncthreads();
  }
  if (threadIdx.x == 0 && grand_total) *grand_total = carry;
}

// phase C: exclusive prefix of every element
template <typename Op>
__global__ void __launch_bounds__(kScanThreads) scan_apply(Op op, long long n, const int* tile_offsets, int* out) {
  const long long base = (long long)blockIdx.x * kScanTile + (long long)threadIdx.x * kScanPer;
  int v[kScanPer], s = 0;
#pragma unroll
  for (int i = 0; i < kScanPer; i++) { v[i] = base + i < n ? op(base + i) : 0; s += v[i]; }
  int ex = block_exclusive_scan(s, nullptr) + tile_offsets[blockIdx.x];
#pragma unroll
  for (int i = 0; i < kScanPer; i++) { if (base + i < n) out[base + i] = ex; ex += v[i]; }
}

template <typename Op>
int exclusive_scan(Op op, long long n, int* out, int* tile_sums, int* grand_total, cudaStream_t st) {
  const int n_tiles = (int)((n + kScanTile - 1) / kScanTile);
  if (n_tiles == 0) return WV_OK;
  scan_tile_sums<<<n_tiles, kScanThreads, 0, st>>>(op, n, tile_sums);
  WV_LAUNCH_CHECK();
  scan_tile_offsets<<<1, kScanThreads, 0, st>>>(tile_sums, n_tiles, grand_total);
  WV_LAUNCH_CHECK();
  scan_apply<<<n_tiles, kScanThreads, 0, st>>>(op, n, tile_sums, out);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

__global__ void mark_keys(const int64_t* __restrict__ keys, long long E, int64_t key_min, int64_t range, uint32_t* bitmap, int* status) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    const int64_t k = keys[e] - key_min;
    if (k < 0 || k >= range) { set_status(status, WV_STATUS_BAD_INDEX); continue; }
    const uint32_t bit = 1u << (k & 31);
    uint32_t* w = bitmap + (k >> 5);
    if (!(*w & bit)) atomicOr(w, bit);   // most keys repeat: skip the atomic once the bit is visible
  }
}

__global__ void rank_keys(const int64_t* __restrict__ keys, long long E, int64_t key_min, int64_t range, const uint32_t* __restrict__ bitmap,
                          const int* __restrict__ word_prefix, int64_t* __restrict__ jx, int* counts) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    const int64_t k = keys[e] - key_min;
    if (k < 0 || k >= range) { jx[e] = 0; continue; }   // flagged by mark_keys; member of no group
    const int g = word_prefix[k >> 5] + __popc(bitmap[k >> 5] & ((1u << (k & 31)) - 1u));
    jx[e] = g;
    atomicAdd(counts + g, 1);
  }
}

__global__ void fill_members(const int64_t* __restrict__ keys, long long E, int64_t key_min, int64_t range, const int64_t* __restrict__ jx,
                             const int* __restrict__ offsets, int* counts, int* members) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
    const int64_t k = keys[e] - key_min;
    if (k < 0 || k >= range) continue;
    const int g = (int)jx[e];
    members[offsets[g] + atomicSub(counts + g, 1) - 1] = (int)e;   // any order inside the group; sorted next
  }
}

// one warp per group: members into ascending edge order (rank sort; groups are a few dozen edges)
__global__ void sort_members(const int* __restrict__ offsets, const int* __restrict__ n_groups, const int* __restrict__ members, int* __restrict__ order) {
  const int lane = threadIdx.x & 31;
  const int G = *n_groups;
  for (int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); g < G; g += gridDim.x * (blockDim.x >> 5)) {
    const int lo = offsets[g], n = offsets[g + 1] - lo;
    if (n <= 32) {
      const int mine = lane < n ? members[lo + lane] : 0x7fffffff;
      int rank = 0;
#pragma unroll
      for (int o = 0; o < 32; o++) rank += __shfl_sync(0xffffffffu, mine, o) < mine ? 1 : 0;
      if (lane < n) order[lo + rank] = mine;
    } else {
      for (int i = lane; i < n; i += 32) {
        const int mine = members[lo + i];
        int rank = 0;
        for (int j = 0; j < n; j++) rank += members[lo + j] < mine ? 1 : 0;
        order[lo + rank] = mine;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// y[b][g][c] = sum_{e in g} f[b][e][c] * exp(s[b][e][c] - max_g) / sum_{e in g} exp(s[b][e][c] - max_g)
// ---------------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float soft_exp(float x);
template <> __device__ __forceinline__ float soft_exp<float>(float x) { return expf(x); }
template <> __device__ __forceinline__ float soft_exp<__half>(float x) { return __expf(x); }   // 2 ulp of fp32: far below the fp16 output rounding

template <typename T, int V> struct Vec;
template <> struct Vec<__half, 8> {
  uint4 r;
  __device__ __forceinline__ void load(const __half* p) { r = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void get(float* o) const {
    const __half2* h = reinterpret_cast<const __half2*>(&r);
#pragma unroll
    for (int i = 0; i < 4; i++) { const float2 f = __half22float2(h[i]); o[2 * i] = f.x; o[2 * i + 1] = f.y; }
  }
  static __device__ __forceinline__ void store(__half* p, const float* v) {
    uint4 w;
    __half2* h = reinterpret_cast<__half2*>(&w);
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = w;
  }
};
template <> struct Vec<__half, 4> {
  uint2 r;
  __device__ __forceinline__ void load(const __half* p) { r = *reinterpret_cast<const uint2*>(p); }
  __device__ __forceinline__ void get(float* o) const {
    const __half2* h = reinterpret_cast<const __half2*>(&r);
#pragma unroll
    for (int i = 0; i < 2; i++) { const float2 f = __half22float2(h[i]); o[2 * i] = f.x; o[2 * i + 1] = f.y; }
  }
  static __device__ __forceinline__ void store(__half* p, const float* v) {
    uint2 w;
    __half2* h = reinterpret_cast<__half2*>(&w);
#pragma unroll
    for (int i = 0; i < 2; i++) h[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint2*>(p) = w;
  }
};
template <> struct Vec<float, 4> {
  float4 r;
  __device__ __forceinline__ void load(const float* p) { r = *reinterpret_cast<const float4*>(p); }
  __device__ __forceinline__ void get(float* o) const { o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = r.w; }
  static __device__ __forceinline__ void store(float* p, const float* v) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }
};
template <typename T> struct Vec<T, 1> {
  T r;
  __device__ __forceinline__ void load(const T* p) { r = *p; }
  __device__ __forceinline__ void get(float* o) const { o[0] = to_f32<T>(r); }
  static __device__ __forceinline__ void store(T* p, const float* v) { *p = from_f32<T>(v[0]); }
};

// A CTA works on GP (batch, group) rows at a time: thread (slot, split, cl) owns V channels and a contiguous share of its group's
// members, which it walks ONCE in chunks of CH: the score rows and value rows of a chunk are loaded together with the member ids of
// the NEXT chunk (2 CH + CH independent loads in flight per thread), the running maximum moves to the chunk's maximum and the partial
// sums are rescaled by exp(old max - new max) (the online form of the same softmax).  With S > 1 the splits' (max, sum, weighted sum)
// triples meet in shared memory and split 0 merges them in split order.  The order of every addition is a function of
// (group size, S, CH) only: deterministic.
template <typename T, int V, int CH>
__global__ void __launch_bounds__(512) soft_agg_kernel(const T* __restrict__ fx, const T* __restrict__ gx, const int* __restrict__ order,
                                                        const int* __restrict__ offsets, const int* __restrict__ n_groups, T* __restrict__ y,
                                                        int B, long long E, int D, long long Gcap, int tpc, int S, int GP) {
  extern __shared__ float red[];                          // [3][GP][S][tpc * V]
  const int cpb = tpc * V;                               // channels per pass of a slot
  const int slot = threadIdx.x / (S * tpc), ts = threadIdx.x - slot * (S * tpc);
  const int split = ts / tpc, cl = ts - split * tpc;
  const size_t plane = (size_t)GP * S * cpb;
  float* r0 = red + (size_t)slot * S * cpb + cl * V;     // split 0's triple of this (slot, channel vector)
  float* rm = r0 + (size_t)split * cpb;                  // this thread's; + plane: sums; + 2 * plane: weighted sums
  const int G = (int)min((long long)*n_groups, Gcap);
  const long long rows = (long long)B * Gcap;
  for (long long bg0 = (long long)blockIdx.x * GP; bg0 < rows; bg0 += (long long)gridDim.x * GP) {
    const long long bg = bg0 + slot;
    const bool live = bg < rows;
    const int g = live ? (int)(bg % Gcap) : 0, b = live ? (int)(bg / Gcap) : 0;
    T* yrow = y + ((long long)b * Gcap + g) * D;
    const bool grp = live && g < G;
    if (live && !grp) {                                   // rows beyond the groups present: defined (zero), never gathered
      for (int c = ts; c < D; c += S * tpc) yrow[c] = from_f32<T>(0.f);
    }
    int m0 = 0, m1 = 0;
    if (grp) {
      const int lo = offsets[g], n = offsets[g + 1] - lo;
      m0 = lo + (int)((long long)n * split / S); m1 = lo + (int)((long long)n * (split + 1) / S);
    }
    for (int c0 = 0; c0 < D; c0 += cpb) {
      const int c = c0 + cl * V;
      const bool act = grp && c < D;
      const T* fb = fx + ((long long)b * E) * D + c;
      const T* gb = gx + ((long long)b * E) * D + c;
      float mx[V], s[V], a[V];
#pragma unroll
      for (int i = 0; i < V; i++) { mx[i] = -INFINITY; s[i] = 0.f; a[i] = 0.f; }
      if (act && m0 < m1) {
        int en[CH];
#pragma unroll
        for (int u = 0; u < CH; u++) en[u] = order[m0 + (m0 + u < m1 ? u : 0)];   // a short chunk repeats its first member (masked below)
        for (int m = m0; m < m1; m += CH) {
          const int cnt = min(CH, m1 - m);
          Vec<T, V> rg[CH], rf[CH];
#pragma unroll
          for (int u = 0; u < CH; u++) { rg[u].load(gb + (long long)en[u] * D); rf[u].load(fb + (long long)en[u] * D); }
          if (m + CH < m1) {
#pragma unroll
            for (int u = 0; u < CH; u++) en[u] = order[m + CH + (m + CH + u < m1 ? u : 0)];
          }
          float nm[V];
#pragma unroll
          for (int i = 0; i < V; i++) nm[i] = mx[i];
#pragma unroll
          for (int u = 0; u < CH; u++) {
            float v[V];
            rg[u].get(v);
#pragma unroll
            for (int i = 0; i < V; i++) nm[i] = fmaxf(nm[i], v[i]);       // (a repeated member cannot raise the maximum)
          }
#pragma unroll
          for (int i = 0; i < V; i++) {
            const float k = soft_exp<T>(mx[i] - nm[i]);                    // first chunk: exp(-inf) = 0 on zero sums
            s[i] *= k; a[i] *= k; mx[i] = nm[i];
          }
#pragma unroll
          for (int u = 0; u < CH; u++) {
            if (u < cnt) {
              float v[V], f[V];
              rg[u].get(v); rf[u].get(f);
#pragma unroll
              for (int i = 0; i < V; i++) { const float w = soft_exp<T>(v[i] - mx[i]); s[i] += w; a[i] = fmaf(f[i], w, a[i]); }
            }
          }
        }
      }
      if (S > 1) {
#pragma unroll
        for (int i = 0; i < V; i++) { rm[i] = mx[i]; rm[plane + i] = s[i]; rm[2 * plane + i] = a[i]; }
        __syncthreads();
        if (split == 0) {
#pragma unroll
          for (int i = 0; i < V; i++) {
            float M = mx[i];
            for (int sp = 1; sp < S; sp++) M = fmaxf(M, r0[(size_t)sp * cpb + i]);
            float ss = 0.f, aa = 0.f;
            for (int sp = 0; sp < S; sp++) {
              const float ms = r0[(size_t)sp * cpb + i];
              if (ms > -INFINITY) {                                        // (a split with no members holds no sum)
                const float k = soft_exp<T>(ms - M);
                ss = fmaf(r0[plane + (size_t)sp * cpb + i], k, ss);
                aa = fmaf(r0[2 * plane + (size_t)sp * cpb + i], k, aa);
              }
            }
            s[i] = ss; a[i] = aa;
          }
        }
      }
      if (split == 0 && act) {
        float o[V];
#pragma unroll
        for (int i = 0; i < V; i++) o[i] = a[i] / s[i];
        Vec<T, V>::store(yrow + c, o);
      }
      if (S > 1) __syncthreads();
    }
  }
}

int agg_env(const char* name, int dflt) {
  const char* s = getenv(name);
  return (s && s[0]) ? atoi(s) : dflt;
}

inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

struct PlanWs {
  uint32_t* bitmap; int* word_prefix; int* tile_sums; int* counts; int* members;
  size_t bytes;
};

PlanWs carve(void* ws, long long E, long long range) {
  const long long W = (range + 31) / 32;
  const long long tiles = (max(W, E + 1) + kScanTile - 1) / kScanTile + 1;
  PlanWs p;
  char* c = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~(uintptr_t)255);
  size_t off = 0;
  p.bitmap = reinterpret_cast<uint32_t*>(c + off); off += al256((size_t)W * 4);
  p.counts = reinterpret_cast<int*>(c + off); off += al256((size_t)(E + 1) * 4);   // adjacent to the bitmap: one memset clears both
  p.word_prefix = reinterpret_cast<int*>(c + off); off += al256((size_t)W * 4);
  p.tile_sums = reinterpret_cast<int*>(c + off); off += al256((size_t)tiles * 4);
  p.members = reinterpret_cast<int*>(c + off); off += al256((size_t)(E + 1) * 4);
  p.bytes = off + 256;
  return p;
}

}  // namespace

// Key ranges up to this go through the counting-sort edge plan shared with the BA (ba_plan.cu: ONE cooperative launch; the
// secondary sort key is the key itself, so a group's members come out in ascending edge order); wider ranges take the bitmap +
// popcount-rank path below (11 small launches, 1 bit per possible key instead of 16 bytes).
constexpr long long kCoopRangeMax = 1LL << 22;

size_t group_plan_workspace_bytes(long long E, long long range) {
  const size_t a = carve(nullptr, E, range).bytes;
  const size_t b = range <= kCoopRangeMax ? plan_bytes(E, range, plan_grid()) + 256 : 0;
  return a > b ? a : b;
}

int group_plan(const int64_t* keys, long long E, int64_t key_min, long long range, int64_t* jx, int* order, int* offsets, int* n_groups,
               void* workspace, size_t workspace_bytes, int* status, cudaStream_t st) {
  if (workspace_bytes < group_plan_workspace_bytes(E, range)) return WV_EWORKSPACE;
  if (range <= kCoopRangeMax && agg_env("WVB200_AGG_BITMAP_PLAN", 0) == 0) {
    Plan pl;
    plan_carve(&pl, reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255), E, range, plan_grid());
    pl.eperm = order;       // the caller's arrays receive the member lists directly
    pl.seg_off = offsets;
    WV_CUDA_TRY(cudaMemsetAsync(jx, 0, (size_t)E * sizeof(int64_t), st));   // (keys outside the range belong to no group)
    return plan_build_groups(pl, keys, key_min, keys, nullptr, 0, nullptr, nullptr, jx, n_groups, status, st);
  }
  const PlanWs p = carve(workspace, E, range);
  const long long W = (range + 31) / 32;
  WV_CUDA_TRY(cudaMemsetAsync(p.bitmap, 0, (size_t)(reinterpret_cast<char*>(p.word_prefix) - reinterpret_cast<char*>(p.bitmap)), st));
  const int tb = 256;
  const int grid = (int)min((long long)sm_count() * 8, (E + tb - 1) / tb);
  mark_keys<<<grid, tb, 0, st>>>(keys, E, key_min, range, p.bitmap, status);
  WV_LAUNCH_CHECK();
  int rc = exclusive_scan(PopcOp{p.bitmap}, W, p.word_prefix, p.tile_sums, n_groups, st);
  if (rc != WV_OK) return rc;
  rank_keys<<<grid, tb, 0, st>>>(keys, E, key_min, range, p.bitmap, p.word_prefix, jx, p.counts);
  WV_LAUNCH_CHECK();
  rc = exclusive_scan(CountOp{p.counts, E}, E + 1, offsets, p.tile_sums, nullptr, st);   // offsets[G .. E] = number of grouped edges
  if (rc != WV_OK) return rc;
  fill_members<<<grid, tb, 0, st>>>(keys, E, key_min, range, jx, offsets, p.counts, p.members);
  WV_LAUNCH_CHECK();
  sort_members<<<(int)min((long long)sm_count() * 8, (E + 7) / 8), 256, 0, st>>>(offsets, n_groups, p.members, order);
  WV_LAUNCH_CHECK();
  return WV_OK;
}


template <typename T, int V>
int launch_soft_agg(const void* fx, const void* gx, const int* order, const int* offsets, const int* n_groups, void* y, int B, long long E,
                    int D, long long Gcap, cudaStream_t st) {
  const int cv = (D + V - 1) / V;                        // channel vectors per row
  const int tpc = cv < 128 ? cv : 128;
  // splits: one thread walks up to ~`per` members of an average group; groups per CTA: fill 256+ threads, keep >= 2 CTAs per SM
  const long long avg = Gcap > 0 ? (E + Gcap - 1) / Gcap : 1;
  // (defaults from a sweep on B200 at E = 45 312, D = 384 fp16, groups of ~21 and ~96 members: profiles/r1_softagg_sweep.txt)
  const int per = agg_env("WVB200_AGG_PER", 48), smax = agg_env("WVB200_AGG_SMAX", 4), ch = agg_env("WVB200_AGG_CHUNK", 4);
  const int tmax = agg_env("WVB200_AGG_THREADS", 96);
  long long S = (avg + per - 1) / per;
  S = S < 1 ? 1 : S;
  S = S > smax ? smax : S;
  while (S > 1 && S * tpc > 512) S--;
  const long long rows = (long long)B * Gcap;
  long long GP = tmax / (S * tpc);
  GP = GP < 1 ? 1 : GP;
  while (GP > 1 && (rows + GP - 1) / GP < 2LL * sm_count()) GP--;
  while ((S > 1 || GP > 1) && S * GP * tpc * V * 12 > 48 * 1024) { if (GP > 1) GP--; else S--; }
  const long long blocks = (rows + GP - 1) / GP;
  const int grid = (int)(blocks < (long long)sm_count() * 64 ? blocks : (long long)sm_count() * 64);
  const size_t smem = S > 1 ? (size_t)S * GP * tpc * V * 12 : 0;
  const int threads = (int)(S * GP * tpc);
  if (ch == 4)
    soft_agg_kernel<T, V, 4><<<grid, threads, smem, st>>>((const T*)fx, (const T*)gx, order, offsets, n_groups, (T*)y, B, E, D, Gcap, tpc, (int)S, (int)GP);
  else
    soft_agg_kernel<T, V, 2><<<grid, threads, smem, st>>>((const T*)fx, (const T*)gx, order, offsets, n_groups, (T*)y, B, E, D, Gcap, tpc, (int)S, (int)GP);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int soft_agg_forward(const void* fx, const void* gx, const int* order, const int* offsets, const int* n_groups, void* y, int dtype,
                     int B, long long E, int D, long long Gcap, cudaStream_t st) {
  const bool al16 = ((reinterpret_cast<uintptr_t>(fx) | reinterpret_cast<uintptr_t>(gx) | reinterpret_cast<uintptr_t>(y)) & 15) == 0;
  if ((long long)B * Gcap == 0) return WV_OK;
  if (dtype == WV_F16) {
    if (D % 8 == 0 && al16 && agg_env("WVB200_AGG_VEC", 8) == 8) return launch_soft_agg<__half, 8>(fx, gx, order, offsets, n_groups, y, B, E, D, Gcap, st);
    if (D % 4 == 0 && al16) return launch_soft_agg<__half, 4>(fx, gx, order, offsets, n_groups, y, B, E, D, Gcap, st);
    return launch_soft_agg<__half, 1>(fx, gx, order, offsets, n_groups, y, B, E, D, Gcap, st);
  }
  if (D % 4 == 0 && al16) return launch_soft_agg<float, 4>(fx, gx, order, offsets, n_groups, y, B, E, D, Gcap, st);
  return launch_soft_agg<float, 1>(fx, gx, order, offsets, n_groups, y, B, E, D, Gcap, st);
}

}  // namespace wv
