// TEST INFRASTRUCTURE ONLY -- not product code.
//
// Host emulation of the handful of CUDA keywords the reference's .cu files use,
// so that the reference's *own sources* (read in place from /root/reference)
// compile with g++ and run on CPU tensors.  Used solely by oracle/build_ref.py
// to build oracle/_ref/*.so, which pins the oracle (oracle/*.c) against the
// reference's arithmetic.  Nothing in the product path includes this file.
//
// What is emulated (no arithmetic is touched):
//   __global__/__device__/__forceinline__  -> plain inline functions
//   __shared__                              -> static storage (threads run serially)
//   __syncthreads()                         -> no-op (serial execution, thread 0 first)
//   blockIdx/threadIdx/blockDim/gridDim     -> globals set by HOST_LAUNCH
//   atomicAdd(p, v)                         -> *p += v  (serial => deterministic order)
//   kernel<<<grid, block>>>(args)           -> HOST_LAUNCH(grid, block, kernel, args)
//                                              (textual rewrite done by build_ref.py)
//   torch::kCUDA                            -> torch::kCPU  (outputs allocated on CPU)
//   torch::linalg::cholesky                 -> at::linalg_cholesky (header dropped in torch>=2.x)
//   AT_DISPATCH(x.type(), ...)              -> overload of ::detail::scalar_type for
//                                              DeprecatedTypeProperties (removed upstream)
#pragma once
#include <torch/extension.h>
#include <ATen/ATen.h>
#include <ATen/NativeFunctions.h>
#include <ATen/Parallel.h>
#include <vector>
#include <iostream>
#include <unordered_map>
#include <algorithm>
#include <cmath>
#include <cstdint>

namespace detail {
inline at::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) {
  return t.scalarType();
}
}  // namespace detail

namespace torch { namespace linalg {
inline at::Tensor cholesky(const at::Tensor& a) { return at::linalg_cholesky(a); }
}}  // namespace torch::linalg

// torch only defines RestrictPtrTraits under __CUDACC__; on the host it is a plain pointer
namespace torch {
template <typename T> struct RestrictPtrTraits { typedef T* PtrType; };
}  // namespace torch

struct host_dim3 { unsigned int x = 0, y = 0, z = 0; };
static host_dim3 blockIdx, threadIdx, blockDim, gridDim;

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __shared__ static
#define __restrict__
inline void __syncthreads() {}

template <typename T, typename U>
inline void atomicAdd(T* p, U v) { *p = *p + static_cast<T>(v); }

// CUDA math overload set picks double for (float, double); mirror that on the host.
inline double max(float a, double b) { return (double)a > b ? (double)a : b; }
inline double max(double a, float b) { return a > (double)b ? a : (double)b; }
inline float max(float a, float b) { return a > b ? a : b; }
using std::floor;
using std::sqrt;

// launch arguments are evaluated once, then the kernel body runs for every
// (block, thread) pair in launch order
#define HOST_LAUNCH(GRID, BLOCK, KERNEL, ...)                              \
  do {                                                                     \
    gridDim.x = (unsigned int)(GRID);                                      \
    blockDim.x = (unsigned int)(BLOCK);                                    \
    auto _run = [&](auto&&... _a) {                                        \
      for (unsigned int _b = 0; _b < gridDim.x; ++_b) {                    \
        blockIdx.x = _b;                                                   \
        for (unsigned int _t = 0; _t < blockDim.x; ++_t) {                 \
          threadIdx.x = _t;                                                \
          KERNEL(_a...);                                                   \
        }                                                                  \
      }                                                                    \
    };                                                                     \
    _run(__VA_ARGS__);                                                     \
  } while (0)

// outputs of the reference are allocated with torch::kCUDA; route them to the CPU
#define kCUDA kCPU
