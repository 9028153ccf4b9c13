// Device-side edge plan (see ba_plan.cuh for the algorithm and the reference lines it replaces).
#include <cooperative_groups.h>
#include <stdlib.h>

#include "ba_plan.cuh"

namespace cg = cooperative_groups;

namespace wv {

static constexpr int kPlanThreads = 512;

static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

size_t plan_bytes(int64_t E, int64_t n_groups, int max_ctas) {
  const int64_t mp_cap = (E < n_groups ? E : n_groups);
  return al256(sizeof(PlanHeader)) + 4 * al256(sizeof(int) * (size_t)n_groups) + 2 * al256(sizeof(int) * (size_t)max_ctas) +
         al256(sizeof(int) * (size_t)(mp_cap + 1)) * 2 + 2 * al256(sizeof(int) * (size_t)(E + 1));
}

size_t plan_carve(Plan* p, void* ws, int64_t E, int64_t n_groups, int max_ctas) {
  char* b = (char*)ws;
  size_t o = 0;
  auto take = [&](size_t bytes) { void* r = b + o; o += al256(bytes); return r; };
  p->E = E; p->n_groups = n_groups; p->Mp_cap = (E < n_groups ? E : n_groups);
  p->hdr = (PlanHeader*)take(sizeof(PlanHeader));
  p->cnt = (int*)take(sizeof(int) * (size_t)n_groups);
  p->cursor = (int*)take(sizeof(int) * (size_t)n_groups);
  p->loc_off = (int*)take(sizeof(int) * (size_t)n_groups);
  p->loc_rank = (int*)take(sizeof(int) * (size_t)n_groups);
  p->chunk_cnt = (int*)take(sizeof(int) * (size_t)max_ctas);
  p->chunk_grp = (int*)take(sizeof(int) * (size_t)max_ctas);
  p->kx = (int*)take(sizeof(int) * (size_t)(p->Mp_cap + 1));
  p->seg_off = (int*)take(sizeof(int) * (size_t)(p->Mp_cap + 1));
  p->tmp = (int*)take(sizeof(int) * (size_t)(E + 1));
  p->eperm = (int*)take(sizeof(int) * (size_t)(E + 1));
  return o;
}

struct PlanArgs {
  Plan p;
  const int64_t* g;
  const int64_t* jj;
  const int64_t* ii;
  int64_t n_poses;
  int64_t* ix;
  int64_t* jx;
  int* status;
  // group-plan form (SoftAgg, agg.cu): keys are g[e] - goff; rank_out[e] = dense rank of e's key (== torch.unique's inverse
  // index), mp_out = number of distinct keys
  int64_t goff;
  int64_t* rank_out;
  int* mp_out;
  // small id spaces (n_groups <= 65 536: the tracker's patch buffer): the host memset that clears the header also clears cnt /
  // cursor for ALL ids, the window is the whole id space, and validation rides along with the count: phases A and B and their
  // two grid syncs disappear
  int small;
};

__device__ __forceinline__ int64_t gkey(const PlanArgs& a, int64_t e) { return a.g[e] - a.goff; }

__device__ __forceinline__ bool edge_valid(const PlanArgs& a, int64_t e) {
  const int64_t gv = gkey(a, e);
  if (gv < 0 || gv >= a.p.n_groups) return false;
  if (a.ii) {
    const int64_t i = a.ii[e], j = a.jj[e];
    if (i < 0 || i >= a.n_poses || j < 0 || j >= a.n_poses) return false;
  }
  return true;
}

// A: window of ids in use + validation
__device__ void plan_phase_minmax(const PlanArgs& a) {
  int mx = 0, mn = 0, bad = 0;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < a.p.E; e += (int64_t)gridDim.x * blockDim.x) {
    if (!edge_valid(a, e)) { bad = 1; continue; }
    const int gv = (int)gkey(a, e);
    mx = max(mx, gv + 1);
    mn = max(mn, (int)a.p.n_groups - gv);
  }
  for (int o = 16; o > 0; o >>= 1) {
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = max(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    bad |= __shfl_xor_sync(0xffffffffu, bad, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (mx) atomicMax(&a.p.hdr->gmax_p1, mx);
    if (mn) atomicMax(&a.p.hdr->gmin_neg, mn);
    if (bad) { atomicOr(&a.p.hdr->bad, 1); set_status(a.status, WV_STATUS_BAD_INDEX); }
  }
}

__device__ __forceinline__ void plan_window(const PlanArgs& a, int& gmin, int& range) {
  if (a.small) { gmin = 0; range = (int)a.p.n_groups; return; }
  const int mxp1 = a.p.hdr->gmax_p1;
  gmin = (int)a.p.n_groups - a.p.hdr->gmin_neg;
  range = mxp1 > 0 ? mxp1 - gmin : 0;
}

// B: zero the counters of the window
__device__ void plan_phase_zero(const PlanArgs& a) {
  int gmin, range; plan_window(a, gmin, range);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < range; i += gridDim.x * blockDim.x) {
    a.p.cnt[i] = 0; a.p.cursor[i] = 0;
  }
}

// C: count
__device__ void plan_phase_count(const PlanArgs& a) {
  int gmin, range; plan_window(a, gmin, range);
  int bad = 0;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < a.p.E; e += (int64_t)gridDim.x * blockDim.x) {
    if (edge_valid(a, e)) atomicAdd(&a.p.cnt[(int)gkey(a, e) - gmin], 1);
    else bad = 1;
  }
  if (a.small && __any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) {   // (otherwise phase A reported it)
    atomicOr(&a.p.hdr->bad, 1); set_status(a.status, WV_STATUS_BAD_INDEX);
  }
}

// block-wide exclusive scan of (v0, v1); returns totals through t0/t1.  blockDim.x == kPlanThreads.
__device__ __forceinline__ void block_scan2(int v0, int v1, int& e0, int& e1, int& t0, int& t1, int* sh /*[2*32+2]*/) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  int s0 = v0, s1 = v1;
  for (int o = 1; o < 32; o <<= 1) {
    const int n0 = __shfl_up_sync(0xffffffffu, s0, o), n1 = __shfl_up_sync(0xffffffffu, s1, o);
    if (lane >= o) { s0 += n0; s1 += n1; }
  }
  if (lane == 31) { sh[w] = s0; sh[32 + w] = s1; }
  __syncthreads();
  if (w == 0) {
    int a0 = lane < nw ? sh[lane] : 0, a1 = lane < nw ? sh[32 + lane] : 0;
    int c0 = a0, c1 = a1;
    for (int o = 1; o < 32; o <<= 1) {
      const int n0 = __shfl_up_sync(0xffffffffu, c0, o), n1 = __shfl_up_sync(0xffffffffu, c1, o);
      if (lane >= o) { c0 += n0; c1 += n1; }
    }
    sh[lane] = c0 - a0; sh[32 + lane] = c1 - a1;
    if (lane == 31) { sh[64] = c0; sh[65] = c1; }
  }
  __syncthreads();
  e0 = sh[w] + s0 - v0; e1 = sh[32 + w] + s1 - v1;
  t0 = sh[64]; t1 = sh[65];
  __syncthreads();
}

// D: chunk-local exclusive scans (CTA c owns window slots [c*chunk, (c+1)*chunk))
__device__ void plan_phase_scan(const PlanArgs& a, int* sh) {
  int gmin, range; plan_window(a, gmin, range);
  const int chunk = (range + gridDim.x - 1) / gridDim.x;
  const int lo = min(range, (int)blockIdx.x * chunk), hi = min(range, lo + chunk);
  int carry0 = 0, carry1 = 0;
  for (int base = lo; base < hi; base += blockDim.x) {
    const int i = base + threadIdx.x;
    const int c = i < hi ? a.p.cnt[i] : 0, f = c > 0 ? 1 : 0;
    int e0, e1, t0, t1;
    block_scan2(c, f, e0, e1, t0, t1, sh);
    if (i < hi) { a.p.loc_off[i] = carry0 + e0; a.p.loc_rank[i] = carry1 + e1; }
    carry0 += t0; carry1 += t1;
  }
  if (threadIdx.x == 0) { a.p.chunk_cnt[blockIdx.x] = carry0; a.p.chunk_grp[blockIdx.x] = carry1; }
}

// E: global bases, materialise kx / seg_off, scatter edges into their segments
__device__ void plan_phase_fill(const PlanArgs& a, int* sh_off /*[gridDim+1]*/, int* sh_rank /*[gridDim+1]*/) {
  int gmin, range; plan_window(a, gmin, range);
  const int chunk = max(1, (range + (int)gridDim.x - 1) / (int)gridDim.x);
  if (threadIdx.x < 32) {  // exclusive scan of the chunk totals by one warp
    int c0 = 0, c1 = 0;
    for (int base = 0; base < (int)gridDim.x; base += 32) {
      const int k = base + threadIdx.x;
      const int v0 = k < (int)gridDim.x ? a.p.chunk_cnt[k] : 0, v1 = k < (int)gridDim.x ? a.p.chunk_grp[k] : 0;
      int s0 = v0, s1 = v1;
      for (int o = 1; o < 32; o <<= 1) {
        const int n0 = __shfl_up_sync(0xffffffffu, s0, o), n1 = __shfl_up_sync(0xffffffffu, s1, o);
        if ((int)threadIdx.x >= o) { s0 += n0; s1 += n1; }
      }
      if (k < (int)gridDim.x) { sh_off[k] = c0 + s0 - v0; sh_rank[k] = c1 + s1 - v1; }
      c0 += __shfl_sync(0xffffffffu, s0, 31); c1 += __shfl_sync(0xffffffffu, s1, 31);
    }
    if (threadIdx.x == 0) { sh_off[gridDim.x] = c0; sh_rank[gridDim.x] = c1; }
  }
  __syncthreads();
  const int lo = min(range, (int)blockIdx.x * chunk), hi = min(range, lo + chunk);
  for (int i = lo + threadIdx.x; i < hi; i += blockDim.x)
    if (a.p.cnt[i] > 0) {
      const int r = sh_rank[blockIdx.x] + a.p.loc_rank[i];
      a.p.kx[r] = gmin + i;
      a.p.seg_off[r] = sh_off[blockIdx.x] + a.p.loc_off[i];
    }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    a.p.hdr->Mp = sh_rank[gridDim.x];
    if (a.mp_out) *a.mp_out = sh_rank[gridDim.x];
    a.p.seg_off[sh_rank[gridDim.x]] = sh_off[gridDim.x];
  }
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < a.p.E; e += (int64_t)gridDim.x * blockDim.x) {
    if (!edge_valid(a, e)) {
      if (a.ix) { a.ix[e] = -1; a.jx[e] = -1; }
      continue;
    }
    const int i = (int)gkey(a, e) - gmin;
    const int off = sh_off[i / chunk] + a.p.loc_off[i];
    const int pos = atomicAdd(&a.p.cursor[i], 1);
    a.p.tmp[off + pos] = (int)e;
  }
}

// F: canonical order inside every segment + neighbors (ba.cpp:128-139)
__device__ void plan_phase_order(const PlanArgs& a) {
  const int Mp = a.p.hdr->Mp;
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int r = warp; r < Mp; r += nwarps) {
    const int off = a.p.seg_off[r], n = a.p.seg_off[r + 1] - off;
    if (n <= 32) {
      const int e = lane < n ? a.p.tmp[off + lane] : 0x7fffffff;
      const long long key = lane < n ? (((long long)a.jj[e] << 32) | (unsigned)e) : 0x7fffffffffffffffLL;
      int rank = 0;
      for (int l = 0; l < n; l++) rank += (__shfl_sync(0xffffffffu, key, l) < key) ? 1 : 0;
      if (lane < n) a.p.eperm[off + rank] = e;
    } else {
      for (int s = lane; s < n; s += 32) {
        const int e = a.p.tmp[off + s];
        const long long key = ((long long)a.jj[e] << 32) | (unsigned)e;
        int rank = 0;
        for (int t = 0; t < n; t++) {
          const int e2 = a.p.tmp[off + t];
          rank += ((((long long)a.jj[e2] << 32) | (unsigned)e2) < key) ? 1 : 0;
        }
        a.p.eperm[off + rank] = e;
      }
    }
    if (a.ix) {
      __syncwarp();
      for (int s = lane; s < n; s += 32) {
        const int e = a.p.eperm[off + s];
        a.ix[e] = s > 0 ? (int64_t)a.p.eperm[off + s - 1] : -1;
        a.jx[e] = s + 1 < n ? (int64_t)a.p.eperm[off + s + 1] : -1;
      }
    }
    if (a.rank_out)
      for (int s = lane; s < n; s += 32) a.rank_out[a.p.tmp[off + s]] = r;
  }
}

// ---- launch forms ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1) plan_coop_kernel(PlanArgs a) {   // 512 threads at tracker sizes, 1024 for > 1 M edges (see plan_build_groups)
  extern __shared__ int sh[];
  cg::grid_group grid = cg::this_grid();
  if (!a.small) {
    plan_phase_minmax(a);
    grid.sync();
    plan_phase_zero(a);
    grid.sync();
  }
  plan_phase_count(a);
  grid.sync();
  plan_phase_scan(a, sh);
  grid.sync();
  plan_phase_fill(a, sh, sh + gridDim.x + 1);
  grid.sync();
  plan_phase_order(a);
}

__global__ void __launch_bounds__(1024, 1) plan_k_minmax(PlanArgs a) { plan_phase_minmax(a); }
__global__ void __launch_bounds__(1024, 1) plan_k_zero(PlanArgs a) { plan_phase_zero(a); }
__global__ void __launch_bounds__(1024, 1) plan_k_count(PlanArgs a) { plan_phase_count(a); }
__global__ void __launch_bounds__(1024, 1) plan_k_scan(PlanArgs a) { extern __shared__ int sh[]; plan_phase_scan(a, sh); }
__global__ void __launch_bounds__(1024, 1) plan_k_fill(PlanArgs a) { extern __shared__ int sh[]; plan_phase_fill(a, sh, sh + gridDim.x + 1); }
__global__ void __launch_bounds__(1024, 1) plan_k_order(PlanArgs a) { plan_phase_order(a); }

bool use_cooperative() {  // read every call so tests can flip it; getenv is a few ns next to a launch
  const char* s = getenv("WVB200_NO_COOP");
  return !(s && s[0] == '1');
}

int plan_grid() { return sm_count(); }

int plan_build(const Plan& p, const int64_t* g, const int64_t* jj, const int64_t* ii, int64_t n_poses, int64_t* ix,
               int64_t* jx, int* status, cudaStream_t st) {
  return plan_build_groups(p, g, 0, jj, ii, n_poses, ix, jx, nullptr, nullptr, status, st);
}

int plan_build_groups(const Plan& p, const int64_t* g, int64_t goff, const int64_t* jj, const int64_t* ii, int64_t n_poses,
                      int64_t* ix, int64_t* jx, int64_t* rank_out, int* mp_out, int* status, cudaStream_t st) {
  PlanArgs a{p, g, jj, ii, n_poses, ix, jx, status, goff, rank_out, mp_out, 0};
  // hdr, cnt and cursor are adjacent in the workspace (plan_carve): one memset clears all three for a small id space
  const char* nosmall = getenv("WVB200_PLAN_WINDOWED");
  a.small = (p.n_groups <= 65536 && !(nosmall && nosmall[0] == '1') &&
             reinterpret_cast<char*>(p.cursor) + sizeof(int) * (size_t)p.n_groups <= reinterpret_cast<char*>(p.loc_off)) ? 1 : 0;
  const size_t clear = a.small ? (size_t)(reinterpret_cast<char*>(p.cursor) - reinterpret_cast<char*>(p.hdr)) + sizeof(int) * (size_t)p.n_groups
                               : sizeof(PlanHeader);
  WV_CUDA_TRY(cudaMemsetAsync(p.hdr, 0, clear, st));
  int grid = plan_grid();   // the workspace is carved for plan_grid() chunks; fewer CTAs are fine (cheaper grid syncs on small edge lists)
  {  // ~700 edges per CTA, at least 32 CTAs: 64 CTAs at the tracker's 45 k edges (31.4 us vs 34.7 us on the full grid, B200 sweep)
    const int64_t want = (p.E + 703) / 704;
    if (want < grid) grid = want < 32 ? 32 : (int)want;
  }
  if (const char* e = getenv("WVB200_PLAN_GRID")) { const int v = atoi(e); if (v > 0 && v <= plan_grid()) grid = v; }
  // at global-BA sizes the edge-striding phases want more threads in flight: 1024 per CTA instead of 512 took the plan of E = 9.8 M
  // edges from 1.09 to 0.77 ms on a B200 (taking four edges per thread and iteration on top of that changed nothing: the phases are
  // bound by the atomics' throughput in L2 by then, not by load latency)
  const int threads = p.E > ((int64_t)1 << 20) ? 1024 : kPlanThreads;
  const size_t sh = sizeof(int) * (size_t)(2 * (grid + 1) + 72);
  if (use_cooperative()) {
    void* args[] = {&a};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)plan_coop_kernel, dim3(grid), dim3(threads), args, sh, st);
    if (e == cudaSuccess) { ++g_launches; return WV_OK; }
    (void)cudaGetLastError();  // fall through to the multi-launch form
  }
  if (!a.small) {
    plan_k_minmax<<<grid, threads, 0, st>>>(a);
    plan_k_zero<<<grid, threads, 0, st>>>(a);
  }
  plan_k_count<<<grid, threads, 0, st>>>(a);
  plan_k_scan<<<grid, threads, sh, st>>>(a);
  plan_k_fill<<<grid, threads, sh, st>>>(a);
  plan_k_order<<<grid, threads, 0, st>>>(a);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

}  // namespace wv
