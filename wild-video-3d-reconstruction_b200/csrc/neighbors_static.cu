// wv_neighbors_static: cuda_ba.neighbors for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
//
// The same device-side edge plan as wv_neighbors (ba_plan.cu), but its scratch -- counters, cursors, segment offsets and the two
// edge permutations -- lives in ONE file-scope `static __device__` array of a fixed, stated size instead of a caller-supplied
// workspace.  Every GPU that loads libwvb200.so holds that array (kScratchBytes = 21 MiB) until the process exits.
//
//   * the array's NAME is a host-side stand-in in host code, not a device address: the address the kernels get comes from
//     cudaGetSymbolAddress, taken once per device with that device current (scratch_for_current_device below);
//   * nothing here calls cudaMalloc / cudaFree or any allocating library routine: plan_build enqueues one cudaMemsetAsync (the
//     plan's header and counters must start at zero; everything else is fully rewritten by every call) and kernel launches;
//   * ALL calls on a device share the array, which keeps the previous call's contents: two calls that overlap on the GPU
//     (different streams, or graph branches without a dependency) race.  Serialise them.
#include "ba_plan.cuh"

namespace wv {

namespace {

constexpr int64_t kMaxEdges = WV_NEIGHBORS_STATIC_MAX_EDGES;     // 262 144: above the tracker's ~191 k active edges at M = 384
constexpr int64_t kMaxGroups = WV_NEIGHBORS_STATIC_MAX_GROUPS;   // 1 048 576 patch ids
constexpr size_t kScratchBytes = (size_t)21 << 20;               // >= plan_bytes(kMaxEdges, kMaxGroups, 148 CTAs) = 20.0 MiB

static __device__ __align__(256) unsigned char g_neighbors_scratch[kScratchBytes];

// device address of the array on the CURRENT device (cached per device; nullptr + *err on failure)
void* scratch_for_current_device(cudaError_t* err) {
  static void* ptrs[64] = {nullptr};
  int dev = 0;
  *err = cudaGetDevice(&dev);
  if (*err != cudaSuccess) return nullptr;
  if (dev >= 0 && dev < 64 && ptrs[dev]) return ptrs[dev];
  void* p = nullptr;
  *err = cudaGetSymbolAddress(&p, g_neighbors_scratch);   // the symbol, not its name as a pointer
  if (*err != cudaSuccess) return nullptr;
  if (dev >= 0 && dev < 64) ptrs[dev] = p;
  return p;
}

}  // namespace
}  // namespace wv

extern "C" int wv_neighbors_static(const int64_t* ii, const int64_t* jj, int64_t E, int64_t n_groups, int64_t* ix, int64_t* jx,
                                   int* status, wv_stream_t stream) {
  using namespace wv;
  if (E < 0 || n_groups <= 0) return WV_EINVAL;
  if (E > kMaxEdges || n_groups > kMaxGroups) return WV_ELIMIT;   // beyond the fixed scratch: use wv_neighbors with a workspace
  if (E == 0) return WV_OK;
  if (!ii || !jj || !ix || !jx) return WV_EINVAL;
  const int G = plan_grid();
  if (plan_bytes(E, n_groups, G) > kScratchBytes) return WV_ELIMIT;
  cudaError_t err = cudaSuccess;
  void* ws = scratch_for_current_device(&err);
  if (!ws) return cuda_fail(err);
  Plan p;
  plan_carve(&p, ws, E, n_groups, G);
  return plan_build(p, ii, jj, nullptr, 0, ix, jx, status, as_stream(stream));
}
