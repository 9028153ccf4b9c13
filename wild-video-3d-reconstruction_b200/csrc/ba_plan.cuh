// Edge plan shared by cuda_ba.forward and cuda_ba.neighbors.
//
// The reference groups edges two ways, both through the host: torch::_unique(kk) with a device sort and a
// host sync for the size (ba_cuda.cu:435-440), and a CPU stable_sort per patch after two D2H copies
// (ba.cpp:106-151).  Here one device-side plan serves both and never touches the host:
//
//   group key g = kk (patch id).  Ids live in a window [gmin, gmax] of the patch buffer, so grouping is a
//   counting sort over that window:  A) min/max + index validation   B) zero the window's counters
//   C) count   D) chunked exclusive scan -> segment offsets and dense ranks (== sorted unique order, so
//   rank == the reference's `ku`, kx[rank] == its `kx`)   E) scatter edge ids into their segment
//   F) one warp per segment orders it by (jj, edge id) -- the stable_sort order of ba.cpp:131-132 -- which
//   makes every later floating-point reduction order-deterministic, and emits `neighbors` for free.
//
// All phases are grid-stride device functions.  ba_plan.cu runs them either as ONE cooperative launch with
// grid syncs between phases or as separate launches.
#pragma once
#include "wv_common.cuh"

namespace wv {

struct PlanHeader {  // device-resident, zeroed (cudaMemsetAsync) at the start of every plan
  int gmax_p1;       // max(g) + 1            (atomicMax from 0)
  int gmin_neg;      // max(n_groups - g)     (atomicMax from 0)  -> gmin = n_groups - gmin_neg
  int Mp;            // number of distinct groups
  int bad;           // != 0: an index was out of range
};

struct Plan {
  PlanHeader* hdr;
  int* cnt;       // [n_groups]   edges per id (window-relative after phase B)
  int* cursor;    // [n_groups]   scatter cursors
  int* loc_off;   // [n_groups]   chunk-local exclusive offsets
  int* loc_rank;  // [n_groups]   chunk-local exclusive ranks
  int* chunk_cnt; // [max_ctas]   per-chunk totals
  int* chunk_grp; // [max_ctas]
  int* kx;        // [Mp_cap]     sorted unique ids                (reference: kx)
  int* seg_off;   // [Mp_cap + 1] segment offsets into eperm
  int* tmp;       // [E]          scatter target (arbitrary order inside a segment)
  int* eperm;     // [E]          final order: by group, then (jj, edge id)
  int64_t E, n_groups, Mp_cap;
};

size_t plan_bytes(int64_t E, int64_t n_groups, int max_ctas);
// carve `p` out of a workspace; returns bytes consumed (256-B aligned pieces)
size_t plan_carve(Plan* p, void* ws, int64_t E, int64_t n_groups, int max_ctas);

// Enqueue the plan for group keys g (=kk), secondary keys jj.  Optional: ii (validated against n_poses),
// ix/jx (neighbors output, may be null).  status receives WV_STATUS_BAD_INDEX.
int plan_build(const Plan& p, const int64_t* g, const int64_t* jj, const int64_t* ii, int64_t n_poses,
               int64_t* ix, int64_t* jx, int* status, cudaStream_t st);

// The same plan for any int64 keys g[e] - goff in [0, n_groups) (SoftAgg's group plan, agg.cu): additionally rank_out[e] = dense
// rank of e's key among the distinct keys (nullable) and *mp_out = their number (nullable).
int plan_build_groups(const Plan& p, const int64_t* g, int64_t goff, const int64_t* jj, const int64_t* ii, int64_t n_poses,
                      int64_t* ix, int64_t* jx, int64_t* rank_out, int* mp_out, int* status, cudaStream_t st);
int plan_grid();

}  // namespace wv
