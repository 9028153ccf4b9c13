// TEST / MEASUREMENT INFRASTRUCTURE ONLY.  Pre-included (nvcc -include) when oracle/build_ref_gpu.py compiles the reference's
// own .cu / .cpp files, where they lie under /root/reference, as real CUDA for sm_100 against torch 2.11.  Two source-compat
// shims, no arithmetic: the reference dispatches on `tensor.type()` (the ::detail::scalar_type overload for
// DeprecatedTypeProperties was removed upstream) and calls torch::linalg::cholesky (header dropped in torch >= 2.x).
#pragma once
#include <torch/extension.h>
#include <ATen/ATen.h>

namespace detail {
inline at::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) { return t.scalarType(); }
}  // namespace detail

namespace torch { namespace linalg {
inline at::Tensor cholesky(const at::Tensor& a) { return at::linalg_cholesky(a); }
}}  // namespace torch::linalg
