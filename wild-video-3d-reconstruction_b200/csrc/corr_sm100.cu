// Tiled sm_100a correlation for the tracker's configuration (placeholder until the tiled kernel lands:
// reports "not supported" so wv_corr_forward takes the generic kernel).
#include "wv_common.cuh"

namespace wv {

bool corr_tiled_supported(int, int, int, int, int, int, int, int, int, int, int) { return false; }
size_t corr_tiled_workspace_bytes(int, int, int, int, int, int, int, int, int, int, int) { return 0; }
int corr_forward_tiled(const void*, const void*, const float*, const int64_t*, const int64_t*, void*, int, int, int, int,
                       int, int, int, int, int, int, int, void*, size_t, int*, cudaStream_t) {
  return WV_EUNSUPPORTED;
}

}  // namespace wv
