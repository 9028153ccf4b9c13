// extern "C" surface of libwvb200.so (declared in include/wvb200.h): argument validation, dispatch between
// the tiled sm_100a kernels and the generic ones, and the host-buffer convenience entries.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include <cstdint>

#include "wv_common.cuh"

namespace wv {
int g_last_cuda_error = 0;
unsigned long long g_launches = 0;

int sm_count() {
  static int per_device[64] = {0};   // one process may drive several GPUs (clip batches in threads): cache per device
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = per_device[dev];
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    per_device[dev] = n;
  }
  return n;
}

// corr_generic.cu
int corr_forward_generic(const void*, const void*, const float*, const int64_t*, const int64_t*, void*, int, int, int, int,
                         int, int, int, int, int, int, int, int*, cudaStream_t);
int corr_backward_generic(const void*, const void*, const float*, const int64_t*, const int64_t*, const float*, float*,
                          float*, int, int, int, int, int, int, int, int, int, int, int, int*, cudaStream_t);
int patchify_forward(const void*, const float*, void*, int, int, int, int, int, int, int, cudaStream_t);
int patchify_backward(const void*, const float*, float*, int, int, int, int, int, int, int, cudaStream_t);
// corr_sm100.cu
bool corr_tiled_supported(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int R);
size_t corr_tiled_workspace_bytes(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int R);
// agg.cu
size_t group_plan_workspace_bytes(long long E, long long range);
int group_plan(const int64_t*, long long, int64_t, long long, int64_t*, int*, int*, int*, void*, size_t, int*, cudaStream_t);
int soft_agg_forward(const void*, const void*, const int*, const int*, const int*, void*, int, int, long long, int, long long, cudaStream_t);
bool corr_stacked_supported(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H0, int W0, int H1, int W1, int R);
size_t corr_stacked_workspace_bytes(int B, int n1, int n2, int H0, int W0, int H1, int W1);
int corr_forward_stacked(const void*, const void*, const void*, const float*, float, float, const int64_t*, const int64_t*, void*, int,
                         int, int, int, int, int, int, int, int, int, int, int, int, void*, size_t, int*, cudaStream_t);
int corr_forward_tiled(const void*, const void*, const float*, const int64_t*, const int64_t*, void*, int, int, int, int,
                       int, int, int, int, int, int, int, void*, size_t, int*, cudaStream_t);
// ba_window.cu
int gridsync_probe(int n_ctas, int iters, cudaStream_t st);
size_t ba_workspace_bytes(int64_t E, int64_t n_patches, int N);
int ba_forward_window(float*, float*, const float*, const float*, const float*, const float*, const int64_t*,
                      const int64_t*, const int64_t*, int64_t, int64_t, int64_t, int, int, int, int, const float*, float,
                      void*, size_t, int*, cudaStream_t, int64_t);
int ba_reproject(const float*, const float*, const float*, const int64_t*, const int64_t*, const int64_t*, int64_t,
                 int64_t, int64_t, int, float*, int*, cudaStream_t);
int ba_neighbors(const int64_t*, const int64_t*, int64_t, int64_t, int64_t*, int64_t*, void*, size_t, int*, cudaStream_t);
size_t plan_bytes(int64_t E, int64_t n_groups, int max_ctas);
int plan_grid();
size_t ba_global_workspace_bytes(int64_t E, int64_t n_patches);
int ba_global_build(float*, float*, const float*, const float*, const float*, const float*, const int64_t*, const int64_t*,
                    const int64_t*, int64_t, int64_t, int64_t, int, int, int, const float*, float, int, int, int, float*, float*,
                    void*, size_t, int*, cudaStream_t, int);
int ba_global_span(const int64_t*, const int64_t*, const int64_t*, int64_t, int64_t, int64_t, int, int, int, int*, void*, size_t, int*, cudaStream_t);
int band_ld(int); size_t band_floats(int, int); int band_max_bs(); size_t band_factor_floats(int, int);
int ba_band_solve(const float*, const float*, int, int, int, float*, float*, int*, cudaStream_t);
int ba_global_damp(float*, int, cudaStream_t);
int ba_global_dz(float*, float*, const float*, const float*, const float*, const float*, const int64_t*, const int64_t*, int64_t,
                 int64_t, int64_t, int, int, int, const float*, float, int, int, const float*, float*, void*, size_t, int*,
                 cudaStream_t);
int ba_global_retract(float*, float*, int64_t, int64_t, int64_t, int, int, int, const float*, const float*, void*, size_t,
                      cudaStream_t);
int pops_jacobian(const float*, const float*, const float*, const int64_t*, const int64_t*, const int64_t*, int64_t, int64_t, int64_t, int, int,
                  float*, float*, float*, float*, int*, cudaStream_t);
int pops_transform(const float*, const float*, const float*, const int64_t*, const int64_t*, const int64_t*, int64_t, int64_t, int64_t,
                   int, int, int, float*, float*, int*, cudaStream_t);
int pops_flow_mag(const float*, const float*, const float*, const int64_t*, const int64_t*, const int64_t*, int64_t, int64_t, int64_t,
                  int, float, float*, int*, cudaStream_t);
int pops_point_cloud(const float*, const float*, const float*, const int64_t*, int64_t, int64_t, int, float*, int*, cudaStream_t);
int frame_distance(const float*, const float*, const float*, int, int, int, float, float*, cudaStream_t);
}  // namespace wv

using namespace wv;

extern "C" {

const char* wv_strerror(int code) {
  switch (code) {
    case WV_OK: return "ok";
    case WV_EINVAL: return "invalid argument";
    case WV_ECUDA: return "CUDA call failed";
    case WV_EWORKSPACE: return "workspace too small";
    case WV_EUNSUPPORTED: return "shape not supported by this build";
    case WV_ELIMIT: return "size beyond the fixed maximum of a no-allocation entry point";
    case 1: return "matrix not positive definite";
    default: return "unknown error";
  }
}

int wv_last_cuda_error(void) { return g_last_cuda_error; }
unsigned long long wv_launch_count(void) { return g_launches; }
int wv_gridsync_probe(int n_ctas, int iters, wv_stream_t stream) { return wv::gridsync_probe(n_ctas, iters, wv::as_stream(stream)); }

int wv_device_supported(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}

// One status word per device, owned by the library, so that every module bound to it (cuda_corr, cuda_ba, cuda_agg,
// lietorch_backends are separate shared objects) reports into -- and reads from -- the same word.
int* wv_status_word(void) {
  static int* words[64] = {nullptr};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  if (!words[dev]) {
    int* p = nullptr;
    if (cudaMalloc(reinterpret_cast<void**>(&p), 256) != cudaSuccess) return nullptr;
    if (cudaMemset(p, 0, 256) != cudaSuccess) { cudaFree(p); return nullptr; }
    words[dev] = p;
  }
  return words[dev];
}

static int word_read(int* status, int n, int* value, wv_stream_t stream, bool clear);

int wv_status_read(int* status, int* value, wv_stream_t stream) { return word_read(status, 1, value, stream, true); }
int wv_words_read(const int* words, int n, int* values, wv_stream_t stream) {
  return word_read(const_cast<int*>(words), n, values, stream, false);
}

// The words travel through a host-MAPPED buffer written by a one-thread kernel that follows the caller's work on the stream, and
// the host spins on a sequence number in that buffer: the result is visible ~4 us after the preceding kernel ends (launch + one
// PCIe write), against ~23 us for cudaMemcpyAsync + event record + event polling (measured on the BA call of bench.py).
__global__ void mirror_words_kernel(const int* __restrict__ src, int n, volatile int* dst, int seq) {
  for (int i = 0; i < n; i++) dst[i] = src[i];
  __threadfence_system();
  dst[16] = seq;
}

static int word_read(int* status, int n, int* value, wv_stream_t stream, bool clear) {
  if (!status || !value || n < 1 || n > 16) return WV_EINVAL;
  // one mapped buffer per device, created on first use (callers hold the GIL / serialise per device)
  struct Chan { volatile int* host = nullptr; int* devp = nullptr; int seq = 0; };
  static Chan chans[64];
  int dev = 0;
  WV_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return WV_EUNSUPPORTED;
  Chan& c = chans[dev];
  if (!c.host) {
    int* h = nullptr;
    WV_CUDA_TRY(cudaHostAlloc(reinterpret_cast<void**>(&h), 32 * sizeof(int), cudaHostAllocMapped));
    for (int i = 0; i < 32; i++) h[i] = 0;
    WV_CUDA_TRY(cudaHostGetDevicePointer(reinterpret_cast<void**>(&c.devp), h, 0));
    c.host = h;
  }
  cudaStream_t st = as_stream(stream);
  const int seq = ++c.seq == 0 ? ++c.seq : c.seq;  // never 0 (the buffer's initial value)
  mirror_words_kernel<<<1, 1, 0, st>>>(status, n, c.devp, seq);
  WV_LAUNCH_CHECK();
  bool seen = false;
  for (long spin = 0; spin < 4000000L; spin++) {        // ~10 ms of polling at most, then a blocking wait (which also surfaces a
    if (c.host[16] == seq) { seen = true; break; }       // device fault of the preceding kernels)
    if ((spin & 0xffff) == 0xffff && cudaStreamQuery(st) != cudaErrorNotReady) break;
  }
  if (!seen) {
    WV_CUDA_TRY(cudaStreamSynchronize(st));
    if (c.host[16] != seq) return cuda_fail(cudaErrorUnknown);
  }
  for (int i = 0; i < n; i++) value[i] = c.host[i];
  if (clear && *value != 0) WV_CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(int), st));
  return WV_OK;
}

static bool corr_args_ok(const void* f1, const void* f2, const float* co, const int64_t* ii, const int64_t* jj, int dtype,
                         int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int R) {
  if (dtype != WV_F16 && dtype != WV_F32) return false;
  if (B < 0 || E < 0 || C <= 0 || PH <= 0 || PW <= 0 || n1 <= 0 || n2 <= 0 || H2 <= 0 || W2 <= 0 || R < 0) return false;
  if ((int64_t)B * E > 0 && (!f1 || !f2 || !co || !ii || !jj)) return false;
  return true;
}

size_t wv_corr_workspace_bytes(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int radius) {
  if (!corr_tiled_supported(dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius)) return 0;
  return corr_tiled_workspace_bytes(dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius);
}

int wv_corr_forward(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii, const int64_t* jj,
                    void* out, int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int radius,
                    void* workspace, size_t workspace_bytes, int* status, wv_stream_t stream) {
  if (!corr_args_ok(fmap1, fmap2, coords, ii, jj, dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius)) return WV_EINVAL;
  if ((int64_t)B * E == 0) return WV_OK;
  if (!out) return WV_EINVAL;
  cudaStream_t st = as_stream(stream);
  if (workspace && corr_tiled_supported(dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius) &&
      workspace_bytes >= corr_tiled_workspace_bytes(dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius)) {
    const int rc = corr_forward_tiled(fmap1, fmap2, coords, ii, jj, out, dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius, workspace,
                                      workspace_bytes, status, st);
    if (rc != WV_EUNSUPPORTED) return rc;  // e.g. a misaligned base pointer: the generic kernel takes any
  }
  return corr_forward_generic(fmap1, fmap2, coords, ii, jj, out, dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius, status, st);
}

size_t wv_group_plan_workspace_bytes(int64_t E, int64_t key_range) {
  if (E < 0 || key_range < 0 || key_range > (1LL << 28) || E > 0x7ffffff0LL) return 0;
  return group_plan_workspace_bytes(E, key_range);
}

int wv_group_plan(const int64_t* keys, int64_t E, int64_t key_min, int64_t key_range, int64_t* jx, int32_t* order, int32_t* offsets,
                  int32_t* n_groups, void* workspace, size_t workspace_bytes, int* status, wv_stream_t stream) {
  if (E < 0 || E > 0x7ffffff0LL || key_range < 0 || key_range > (1LL << 28) || !offsets || !n_groups || !workspace) return WV_EINVAL;
  if (E > 0 && (!keys || !jx || !order)) return WV_EINVAL;
  if (E > 0 && key_range == 0) return WV_EINVAL;
  return group_plan(keys, E, key_min, key_range, jx, order, offsets, n_groups, workspace, workspace_bytes, status, as_stream(stream));
}

int wv_soft_agg_forward(const void* fx, const void* gx, const int32_t* order, const int32_t* offsets, const int32_t* n_groups, void* y,
                        int dtype, int B, int64_t E, int D, int64_t g_capacity, wv_stream_t stream) {
  if ((dtype != WV_F16 && dtype != WV_F32) || B < 0 || E < 0 || D <= 0 || g_capacity < 0 || g_capacity > 0x7fffffffLL) return WV_EINVAL;
  if ((int64_t)B * g_capacity == 0) return WV_OK;
  if (!fx || !gx || !order || !offsets || !n_groups || !y) return WV_EINVAL;
  return soft_agg_forward(fx, gx, order, offsets, n_groups, y, dtype, B, E, D, g_capacity, as_stream(stream));
}

size_t wv_corr_stacked_workspace_bytes(int dtype, int B, int E, int C, int PH, int PW, int n1, int n2, int H0, int W0, int H1,
                                       int W1, int radius) {
  if (!corr_stacked_supported(dtype, B, E, C, PH, PW, n1, n2, H0, W0, H1, W1, radius)) return 0;
  return corr_stacked_workspace_bytes(B, n1, n2, H0, W0, H1, W1);
}

int wv_corr_forward_stacked(const void* fmap1, const void* fmap2_l0, const void* fmap2_l1, const float* coords, float cmul0,
                            float cmul1, const int64_t* ii, const int64_t* jj, void* out, int dtype, int B, int E, int C, int PH,
                            int PW, int n1, int n2, int H0, int W0, int H1, int W1, int radius, void* workspace,
                            size_t workspace_bytes, int* status, wv_stream_t stream) {
  if (!corr_args_ok(fmap1, fmap2_l0, coords, ii, jj, dtype, B, E, C, PH, PW, n1, n2, H0, W0, radius) || H1 <= 0 || W1 <= 0)
    return WV_EINVAL;
  if (!(cmul0 == cmul0) || !(cmul1 == cmul1)) return WV_EINVAL;
  if ((int64_t)B * E == 0) return WV_OK;
  if (!out || !fmap2_l1 || !workspace) return WV_EINVAL;
  return corr_forward_stacked(fmap1, fmap2_l0, fmap2_l1, coords, cmul0, cmul1, ii, jj, out, dtype, B, E, C, PH, PW, n1, n2, H0, W0,
                              H1, W1, radius, workspace, workspace_bytes, status, as_stream(stream));
}

int wv_corr_backward(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii, const int64_t* jj,
                     const float* grad, float* fmap1_grad, float* fmap2_grad, int dtype, int B, int E, int C, int PH,
                     int PW, int n1, int n2, int H2, int W2, int radius, int* status, wv_stream_t stream) {
  if (!corr_args_ok(fmap1, fmap2, coords, ii, jj, dtype, B, E, C, PH, PW, n1, n2, H2, W2, radius)) return WV_EINVAL;
  if ((int64_t)B * E == 0) return WV_OK;
  if (!grad || !fmap1_grad || !fmap2_grad) return WV_EINVAL;
  return corr_backward_generic(fmap1, fmap2, coords, ii, jj, grad, fmap1_grad, fmap2_grad, dtype, B, E, C, PH, PW, n1, n2,
                               H2, W2, radius, status, as_stream(stream));
}

int wv_patchify_forward(const void* net, const float* coords, void* patches, int dtype, int B, int M, int C, int H, int W,
                        int radius, wv_stream_t stream) {
  if ((dtype != WV_F16 && dtype != WV_F32) || B < 0 || M < 0 || C <= 0 || H <= 0 || W <= 0 || radius < 0) return WV_EINVAL;
  if ((int64_t)B * M == 0) return WV_OK;
  if (!net || !coords || !patches) return WV_EINVAL;
  return patchify_forward(net, coords, patches, dtype, B, M, C, H, W, radius, as_stream(stream));
}

int wv_patchify_backward(const void* gradient, const float* coords, float* net_grad, int dtype, int B, int M, int C, int H,
                         int W, int radius, wv_stream_t stream) {
  if ((dtype != WV_F16 && dtype != WV_F32) || B < 0 || M < 0 || C <= 0 || H <= 0 || W <= 0 || radius < 0) return WV_EINVAL;
  if ((int64_t)B * M == 0) return WV_OK;
  if (!gradient || !coords || !net_grad) return WV_EINVAL;
  return patchify_backward(gradient, coords, net_grad, dtype, B, M, C, H, W, radius, as_stream(stream));
}

size_t wv_ba_workspace_bytes(int64_t E, int64_t n_patches, int n_free_poses) {
  if (E < 0 || n_patches < 0 || n_free_poses < 0) return 0;
  return ba_workspace_bytes(E, n_patches, n_free_poses);
}

int wv_ba_forward(float* poses, float* patches, const float* intrinsics, const float* target, const float* weight,
                  const float* lmbda, const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses,
                  int64_t n_patches, int P, int t0, int t1, int iterations, const float* patches_est, float mu,
                  void* workspace, size_t workspace_bytes, int* status, wv_stream_t stream) {
  if (E < 0 || n_poses <= 0 || n_patches <= 0 || P <= 0 || t0 < 0 || t1 < t0 || t1 > n_poses) return WV_EINVAL;
  if (E >= (int64_t)1 << 31 || n_patches >= (int64_t)1 << 31) return WV_EUNSUPPORTED;
  if (E == 0 || iterations <= 0) return WV_OK;
  if (!poses || !patches || !intrinsics || !target || !weight || !lmbda || !ii || !jj || !kk || !workspace) return WV_EINVAL;
  return ba_forward_window(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, E, n_poses, n_patches, P, t0, t1,
                           iterations, patches_est, mu, workspace, workspace_bytes, status, as_stream(stream), 0);
}

int wv_ba_forward_planned(float* poses, float* patches, const float* intrinsics, const float* target, const float* weight,
                          const float* lmbda, const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                          int64_t n_poses, int64_t n_patches, int P, int t0, int t1, int iterations, const float* patches_est,
                          float mu, void* workspace, size_t workspace_bytes, int64_t plan_groups, int* status,
                          wv_stream_t stream) {
  if (E < 0 || n_poses <= 0 || n_patches <= 0 || P <= 0 || t0 < 0 || t1 < t0 || t1 > n_poses || plan_groups <= 0) return WV_EINVAL;
  if (E >= (int64_t)1 << 31 || n_patches > plan_groups) return WV_EUNSUPPORTED;
  if (E == 0 || iterations <= 0) return WV_OK;
  if (!poses || !patches || !intrinsics || !target || !weight || !lmbda || !ii || !jj || !kk || !workspace) return WV_EINVAL;
  return ba_forward_window(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, E, n_poses, n_patches, P, t0, t1,
                           iterations, patches_est, mu, workspace, workspace_bytes, status, as_stream(stream), plan_groups);
}

size_t wv_ba_global_workspace_bytes(int64_t E, int64_t n_patches) {
  if (E < 0 || n_patches < 0) return 0;
  return ba_global_workspace_bytes(E, n_patches);
}

static int global_args_ok(int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0, int t1, int shard, int nshards) {
  if (E < 0 || n_poses <= 0 || n_patches <= 0 || P <= 0 || t0 < 0 || t1 < t0 || t1 > n_poses || nshards <= 0 || shard < 0 ||
      shard >= nshards)
    return WV_EINVAL;
  if (E >= (int64_t)1 << 31 || n_patches >= (int64_t)1 << 31 || (int64_t)(t1 - t0) * 6 >= 46340) return WV_EUNSUPPORTED;
  return WV_OK;
}

int wv_ba_global_build(float* poses, float* patches, const float* intrinsics, const float* target, const float* weight,
                       const float* lmbda, const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses,
                       int64_t n_patches, int P, int t0, int t1, const float* patches_est, float mu, int shard, int nshards,
                       int build_plan, float* S, float* y, void* workspace, size_t workspace_bytes, int* status,
                       wv_stream_t stream) {
  const int rc = global_args_ok(E, n_poses, n_patches, P, t0, t1, shard, nshards);
  if (rc != WV_OK) return rc;
  if (!poses || !patches || !intrinsics || !target || !weight || !lmbda || !ii || !jj || !kk || !workspace) return WV_EINVAL;
  if (t1 > t0 && (!S || !y)) return WV_EINVAL;
  if ((reinterpret_cast<uintptr_t>(S) & 15) != 0) return WV_EINVAL;  // vector REDs into S
  return ba_global_build(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, E, n_poses, n_patches, P, t0, t1,
                         patches_est, mu, shard, nshards, build_plan, S, y, workspace, workspace_bytes, status, as_stream(stream), 0);
}

int wv_ba_band_max_bs(void) { return band_max_bs(); }
size_t wv_ba_band_floats(int n6, int bs) { return band_floats(n6, bs); }
size_t wv_ba_band_factor_floats(int n6, int bs) { return band_factor_floats(n6, bs); }

int wv_ba_global_span(const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int t0,
                      int t1, int build_plan, int* span_out, void* workspace, size_t workspace_bytes, int* status, wv_stream_t stream) {
  if (E <= 0 || n_poses <= 0 || n_patches <= 0 || !ii || !jj || !kk || !span_out || !workspace || t1 < t0) return WV_EINVAL;
  if (E >= (int64_t)1 << 31 || n_patches >= (int64_t)1 << 31) return WV_EUNSUPPORTED;
  return ba_global_span(ii, jj, kk, E, n_poses, n_patches, t0, t1, build_plan, span_out, workspace, workspace_bytes, status, as_stream(stream));
}

int wv_ba_global_build_band(float* poses, float* patches, const float* intrinsics, const float* target, const float* weight,
                            const float* lmbda, const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E, int64_t n_poses,
                            int64_t n_patches, int P, int t0, int t1, const float* patches_est, float mu, int shard, int nshards,
                            int build_plan, int bs, float* Sband, float* y, void* workspace, size_t workspace_bytes, int* status,
                            wv_stream_t stream) {
  const int rc = global_args_ok(E, n_poses, n_patches, P, t0, t1, shard, nshards);
  if (rc != WV_OK) return rc;
  if (!poses || !patches || !intrinsics || !target || !weight || !lmbda || !ii || !jj || !kk || !workspace || !Sband || !y) return WV_EINVAL;
  if (bs < 2 || bs > band_max_bs() || (reinterpret_cast<uintptr_t>(Sband) & 15) != 0) return WV_EINVAL;
  return ba_global_build(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, E, n_poses, n_patches, P, t0, t1,
                         patches_est, mu, shard, nshards, build_plan, Sband, y, workspace, workspace_bytes, status, as_stream(stream), bs);
}

int wv_ba_band_solve(const float* Sband, const float* y, int n6, int bs, int damp, float* factors, float* dX, int* status,
                     wv_stream_t stream) {
  if (n6 < 0 || (n6 > 0 && (!Sband || !y || !factors || !dX))) return WV_EINVAL;
  if ((reinterpret_cast<uintptr_t>(factors) & 15) != 0 || (reinterpret_cast<uintptr_t>(Sband) & 15) != 0) return WV_EINVAL;
  return ba_band_solve(Sband, y, n6, bs, damp, factors, dX, status, as_stream(stream));
}

int wv_ba_global_damp(float* S, int n6, wv_stream_t stream) {
  if (n6 < 0 || (n6 > 0 && !S)) return WV_EINVAL;
  return ba_global_damp(S, n6, as_stream(stream));
}

int wv_ba_global_dz(float* poses, float* patches, const float* intrinsics, const float* target, const float* weight,
                    const float* lmbda, const int64_t* ii, const int64_t* jj, int64_t E, int64_t n_poses, int64_t n_patches, int P,
                    int t0, int t1, const float* patches_est, float mu, int shard, int nshards, const float* dX, float* dZ,
                    void* workspace, size_t workspace_bytes, int* status, wv_stream_t stream) {
  const int rc = global_args_ok(E, n_poses, n_patches, P, t0, t1, shard, nshards);
  if (rc != WV_OK) return rc;
  if (!poses || !patches || !intrinsics || !target || !weight || !lmbda || !ii || !jj || !dZ || !workspace) return WV_EINVAL;
  if (t1 > t0 && !dX) return WV_EINVAL;
  return ba_global_dz(poses, patches, intrinsics, target, weight, lmbda, ii, jj, E, n_poses, n_patches, P, t0, t1, patches_est, mu,
                      shard, nshards, dX, dZ, workspace, workspace_bytes, status, as_stream(stream));
}

int wv_ba_global_retract(float* poses, float* patches, int64_t E, int64_t n_poses, int64_t n_patches, int P, int t0, int t1,
                         const float* dX, const float* dZ, void* workspace, size_t workspace_bytes, wv_stream_t stream) {
  const int rc = global_args_ok(E, n_poses, n_patches, P, t0, t1, 0, 1);
  if (rc != WV_OK) return rc;
  if (!poses || !patches || !dZ || !workspace || (t1 > t0 && !dX)) return WV_EINVAL;
  return ba_global_retract(poses, patches, E, n_poses, n_patches, P, t0, t1, dX, dZ, workspace, workspace_bytes, as_stream(stream));
}

int wv_ba_reproject(const float* poses, const float* patches, const float* intrinsics, const int64_t* ii, const int64_t* jj,
                    const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, float* coords, int* status,
                    wv_stream_t stream) {
  if (E < 0 || n_poses <= 0 || n_patches <= 0 || P <= 0) return WV_EINVAL;
  if (E == 0) return WV_OK;
  if (!poses || !patches || !intrinsics || !ii || !jj || !kk || !coords) return WV_EINVAL;
  return ba_reproject(poses, patches, intrinsics, ii, jj, kk, E, n_poses, n_patches, P, coords, status, as_stream(stream));
}

int wv_pops_transform(const float* poses, const float* patches, const float* intrinsics, const int64_t* ii, const int64_t* jj,
                      const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, int tonly, int depth, float* coords,
                      float* valid, int* status, wv_stream_t stream) {
  if (E < 0 || n_poses <= 0 || n_patches <= 0 || P <= 0) return WV_EINVAL;
  if (E == 0) return WV_OK;
  if (!poses || !patches || !intrinsics || !ii || !jj || !kk || !coords) return WV_EINVAL;
  return pops_transform(poses, patches, intrinsics, ii, jj, kk, E, n_poses, n_patches, P, tonly != 0, depth != 0, coords, valid, status,
                        as_stream(stream));
}

int wv_pops_jacobian(const float* poses, const float* patches, const float* intrinsics, const int64_t* ii, const int64_t* jj,
                     const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, int tonly, float* Ji, float* Jj, float* Jz,
                     float* validc, int* status, wv_stream_t stream) {
  if (E < 0 || n_poses <= 0 || n_patches <= 0 || P <= 0) return WV_EINVAL;
  if (E == 0) return WV_OK;
  if (!poses || !patches || !intrinsics || !ii || !jj || !kk || !Ji || !Jj || !Jz || !validc) return WV_EINVAL;
  return pops_jacobian(poses, patches, intrinsics, ii, jj, kk, E, n_poses, n_patches, P, tonly != 0, Ji, Jj, Jz, validc, status, as_stream(stream));
}

int wv_pops_flow_mag(const float* poses, const float* patches, const float* intrinsics, const int64_t* ii, const int64_t* jj,
                     const int64_t* kk, int64_t E, int64_t n_poses, int64_t n_patches, int P, float beta, float* out, int* status,
                     wv_stream_t stream) {
  if (E < 0 || n_poses <= 0 || n_patches <= 0 || P <= 0) return WV_EINVAL;
  if (E == 0) return WV_OK;
  if (!poses || !patches || !intrinsics || !ii || !jj || !kk || !out) return WV_EINVAL;
  return pops_flow_mag(poses, patches, intrinsics, ii, jj, kk, E, n_poses, n_patches, P, beta, out, status, as_stream(stream));
}

int wv_pops_point_cloud(const float* poses, const float* patches, const float* intrinsics, const int64_t* ix, int64_t n,
                        int64_t n_poses, int P, float* out, int* status, wv_stream_t stream) {
  if (n < 0 || n_poses <= 0 || P <= 0) return WV_EINVAL;
  if (n == 0) return WV_OK;
  if (!poses || !patches || !intrinsics || !ix || !out) return WV_EINVAL;
  return pops_point_cloud(poses, patches, intrinsics, ix, n, n_poses, P, out, status, as_stream(stream));
}

int wv_frame_distance(const float* poses, const float* patches, const float* intrinsics, int64_t n, int64_t n_poses,
                      int64_t n_patches, int M, int P, float beta, float* D, wv_stream_t stream) {
  if (n < 0 || M <= 0 || P <= 0 || n > n_poses || n * M > n_patches) return WV_EINVAL;
  if (n > 65535) return WV_EUNSUPPORTED;
  if (n == 0) return WV_OK;
  if (!poses || !patches || !intrinsics || !D) return WV_EINVAL;
  return frame_distance(poses, patches, intrinsics, (int)n, M, P, beta, D, as_stream(stream));
}

int wv_neighbors(const int64_t* ii, const int64_t* jj, int64_t E, int64_t n_groups, int64_t* ix, int64_t* jx,
                 void* workspace, size_t workspace_bytes, int* status, wv_stream_t stream) {
  if (E < 0 || n_groups <= 0) return WV_EINVAL;
  if (E >= (int64_t)1 << 31 || n_groups >= (int64_t)1 << 31) return WV_EUNSUPPORTED;
  if (E == 0) return WV_OK;
  if (!ii || !jj || !ix || !jx || !workspace) return WV_EINVAL;
  return ba_neighbors(ii, jj, E, n_groups, ix, jx, workspace, workspace_bytes, status, as_stream(stream));
}

}  // extern "C"
