// Shared helpers for the sm_100a kernels behind include/wvb200.h.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/wvb200.h"

namespace wv {

extern int g_last_cuda_error;  // abi.cu
extern unsigned long long g_launches;  // abi.cu: kernels this library has launched (wv_launch_count)

inline int cuda_fail(cudaError_t e) {
  g_last_cuda_error = (int)e;
  return WV_ECUDA;
}

#define WV_CUDA_TRY(expr)                                   \
  do {                                                      \
    cudaError_t _e = (expr);                                \
    if (_e != cudaSuccess) return ::wv::cuda_fail(_e);      \
  } while (0)

#define WV_LAUNCH_CHECK()               \
  do {                                  \
    ++::wv::g_launches;                 \
    WV_CUDA_TRY(cudaGetLastError());    \
  } while (0)

inline cudaStream_t as_stream(wv_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// number of SMs of the current device (148 on B200), cached
int sm_count();

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

__device__ __forceinline__ void set_status(int* status, int bit) {
  if (status) atomicOr(status, bit);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace wv
