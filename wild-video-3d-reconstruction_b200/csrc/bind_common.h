// Shared glue of the two torch extension modules: device guard, current stream, a persistent per-device
// status word, error-code -> exception translation.  No kernels here: everything goes through the C ABI.
#pragma once
#include <torch/extension.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <cstdlib>
#include <map>
#include <vector>

#include "wvb200.h"

namespace wvbind {

inline bool env_flag(const char* name) {
  const char* s = std::getenv(name);
  return s && s[0] == '1';
}

inline wv_stream_t current_stream() {
  return reinterpret_cast<wv_stream_t>(c10::cuda::getCurrentCUDAStream().stream());
}

inline void check_cuda(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor");
}

inline void check_rc(int rc, const char* what) {
  if (rc == WV_OK) return;
  if (rc == WV_ECUDA) {
    TORCH_CHECK(false, what, ": CUDA error ", wv_last_cuda_error(), " (", wv_strerror(rc), ")");
  }
  TORCH_CHECK(false, what, ": ", wv_strerror(rc));
}

// the library's int32 status word of this device (wv_status_word: shared by all extension modules), zero when healthy
inline torch::Tensor& status_word(const torch::Device& dev) {
  static std::map<int, torch::Tensor> words;
  auto it = words.find(dev.index());
  if (it == words.end()) {
    c10::cuda::CUDAGuard guard(dev);
    int* p = wv_status_word();
    TORCH_CHECK(p != nullptr, "wv_status_word: could not allocate the status word on ", dev);
    it = words.emplace(dev.index(), torch::from_blob(p, {1}, torch::dtype(torch::kInt32).device(dev))).first;
  }
  return it->second;
}

// read + clear the status word (host sync)
inline int read_status(const torch::Device& dev, const char* what) {
  torch::Tensor& w = status_word(dev);
  int s = 0;
  check_rc(wv_status_read(w.data_ptr<int>(), &s, current_stream()), what);  // async copy to a pinned word + event spin; clears when set
  return s;
}

// raise like the reference would
inline void raise_for_status(int s, const char* what) {
  if (s == 0) return;
  TORCH_CHECK_INDEX(!(s & WV_STATUS_BAD_INDEX), what, ": index out of range in the edge list");
  TORCH_CHECK(!(s & WV_STATUS_STALE_PLAN), what, ": the cached edge plan did not match kk and the call ran with WVB200_ASYNC=1, so it could "
              "not be repeated with a fresh plan (poses / patches were left untouched); set WVB200_NO_PLAN_CACHE=1");
  TORCH_CHECK(!(s & WV_STATUS_TOO_DENSE), what, ": a patch has more than 8256 pose entries (global BA entry list: duplicate edges?)");
  TORCH_CHECK(!(s & WV_STATUS_NOT_SPD), what,
              ": linalg.cholesky: the reduced camera system is not positive-definite");
}

inline void raise_on_status(const torch::Device& dev, const char* what) { raise_for_status(read_status(dev, what), what); }

inline int wv_dtype_of(const torch::Tensor& t, const char* name) {
  if (t.scalar_type() == torch::kHalf) return WV_F16;
  if (t.scalar_type() == torch::kFloat) return WV_F32;
  TORCH_CHECK(false, name, ": expected a half or float tensor, got ", t.scalar_type());
  return -1;
}

}  // namespace wvbind
