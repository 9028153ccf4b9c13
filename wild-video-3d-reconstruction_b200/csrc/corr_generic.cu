// Generic-shape altcorr kernels (any C, patch size, radius; fp16 or fp32 maps).
//
// These cover every shape the reference accepts.  The tracker's own configuration (fp16 maps, C = 128,
// 3x3 patches, radius 3) is routed to the tiled kernel in corr_sm100.cu by wv_corr_forward; this file is
// what runs for training-time fp32 maps, other radii, patchify and the backward passes.
//
// Semantics follow dpvo/altcorr/correlation_kernel.cu (cited per kernel); the implementation does not:
// one CTA per edge, the 3x3xC source patch staged in shared memory as fp32, taps laid out so that a warp
// reads consecutive x positions of one channel plane, fp32 accumulation, and the 4-corner bilinear blend
// + output permutation fused into the same kernel (the reference runs ~9 ATen kernels over a [E,8,8,3,3]
// temporary for that).
#include "wv_common.cuh"

namespace wv {

// ---------------------------------------------------------------------------------------------------
// forward: correlation_kernel.cu:82-136 (raw taps) + :221-232 (blend, permute)
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128) corr_fwd_generic_kernel(
    const T* __restrict__ fmap1, const T* __restrict__ fmap2, const float* __restrict__ coords,
    const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, T* __restrict__ out, int E, int C,
    int PH, int PW, int n1, int n2, int H2, int W2, int R, int* status) {
  extern __shared__ float smem[];
  const int PP = PH * PW, D = 2 * R + 2, TT = D - 1;
  float* f1s = smem;             // [C][PP]
  float* raw = f1s + C * PP;     // [PP][D][D]
  float* cxy = raw + PP * D * D; // [2][PP]
  const int be = blockIdx.x, n = be / E, e = be % E;
  const int64_t ix = ii[e], jx = jj[e];
  T* o = out + (size_t)be * TT * TT * PP;
  if (ix < 0 || ix >= n1 || jx < 0 || jx >= n2) {
    if (threadIdx.x == 0) set_status(status, WV_STATUS_BAD_INDEX);
    for (int t = threadIdx.x; t < TT * TT * PP; t += blockDim.x) o[t] = from_f32<T>(0.f);
    return;
  }
  const T* g1 = fmap1 + ((size_t)n * n1 + ix) * C * PP;
  const T* g2 = fmap2 + ((size_t)n * n2 + jx) * C * H2 * W2;
  for (int t = threadIdx.x; t < C * PP; t += blockDim.x) f1s[t] = to_f32(g1[t]);
  for (int t = threadIdx.x; t < 2 * PP; t += blockDim.x) cxy[t] = coords[(size_t)be * 2 * PP + t];
  __syncthreads();

  const size_t plane = (size_t)H2 * W2;
  for (int t = threadIdx.x; t < PP * D * D; t += blockDim.x) {
    const int b = t % D, a = (t / D) % D, p = t / (D * D);
    const int i1 = (int)floorf(cxy[PP + p]) + (a - R);
    const int j1 = (int)floorf(cxy[p]) + (b - R);
    float s = 0.f;
    if (i1 >= 0 && i1 < H2 && j1 >= 0 && j1 < W2) {
      const T* p2 = g2 + (size_t)i1 * W2 + j1;
      int c = 0;
      for (; c + 8 <= C; c += 8) {
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; q++) v[q] = to_f32(p2[(size_t)(c + q) * plane]);
#pragma unroll
        for (int q = 0; q < 8; q++) s = fmaf(f1s[(c + q) * PP + p], v[q], s);
      }
      for (; c < C; c++) s = fmaf(f1s[c * PP + p], to_f32(p2[(size_t)c * plane]), s);
    }
    raw[t] = s;
  }
  __syncthreads();

  // out[x-tap b][y-tap a][i0][j0]; consecutive threads write consecutive elements
  for (int t = threadIdx.x; t < TT * TT * PP; t += blockDim.x) {
    const int p = t % PP, a = (t / PP) % TT, b = t / (PP * TT);
    const float x = cxy[p], y = cxy[PP + p];
    const float dx = x - floorf(x), dy = y - floorf(y);
    const float* r = raw + (p * D + a) * D + b;
    float v = ((1.f - dx) * (1.f - dy)) * r[0];
    v += (dx * (1.f - dy)) * r[1];
    v += ((1.f - dx) * dy) * r[D];
    v += (dx * dy) * r[D + 1];
    o[t] = from_f32<T>(v);
  }
}

template <typename T>
static int launch_corr_fwd_generic(const void* fmap1, const void* fmap2, const float* coords,
                                   const int64_t* ii, const int64_t* jj, void* out, int B, int E, int C,
                                   int PH, int PW, int n1, int n2, int H2, int W2, int R, int* status,
                                   cudaStream_t st) {
  const int PP = PH * PW, D = 2 * R + 2;
  const size_t smem = sizeof(float) * ((size_t)C * PP + (size_t)PP * D * D + 2 * PP);
  if (smem > 200 * 1024) return WV_EUNSUPPORTED;
  if (smem > 48 * 1024)
    WV_CUDA_TRY(cudaFuncSetAttribute(corr_fwd_generic_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  corr_fwd_generic_kernel<T><<<B * E, 128, smem, st>>>((const T*)fmap1, (const T*)fmap2, coords, ii, jj, (T*)out,
                                                      E, C, PH, PW, n1, n2, H2, W2, R, status);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int corr_forward_generic(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii,
                         const int64_t* jj, void* out, int dtype, int B, int E, int C, int PH, int PW, int n1,
                         int n2, int H2, int W2, int R, int* status, cudaStream_t st) {
  if (dtype == WV_F16)
    return launch_corr_fwd_generic<__half>(fmap1, fmap2, coords, ii, jj, out, B, E, C, PH, PW, n1, n2, H2, W2, R, status, st);
  return launch_corr_fwd_generic<float>(fmap1, fmap2, coords, ii, jj, out, B, E, C, PH, PW, n1, n2, H2, W2, R, status, st);
}

// ---------------------------------------------------------------------------------------------------
// backward: correlation_kernel.cu:139-190 (scatter) + :252-271 (spread of the blended gradient)
// fp32 accumulators (the reference accumulates atomically in scalar_t, i.e. in half for fp16 maps)
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128) corr_bwd_generic_kernel(
    const T* __restrict__ fmap1, const T* __restrict__ fmap2, const float* __restrict__ coords,
    const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, const float* __restrict__ grad,
    float* __restrict__ fmap1_grad, float* __restrict__ fmap2_grad, int E, int C, int PH, int PW, int n1,
    int n2, int H2, int W2, int R, int* status) {
  extern __shared__ float smem[];
  const int PP = PH * PW, D = 2 * R + 2, TT = D - 1;
  float* f1s = smem;             // [C][PP]
  float* cg = f1s + C * PP;      // [PP][D][D]
  float* cxy = cg + PP * D * D;  // [2][PP]
  const int be = blockIdx.x, n = be / E, e = be % E;
  const int64_t ix = ii[e], jx = jj[e];
  if (ix < 0 || ix >= n1 || jx < 0 || jx >= n2) {
    if (threadIdx.x == 0) set_status(status, WV_STATUS_BAD_INDEX);
    return;
  }
  const size_t o1 = ((size_t)n * n1 + ix) * C * PP, o2 = ((size_t)n * n2 + jx) * C * H2 * W2;
  for (int t = threadIdx.x; t < C * PP; t += blockDim.x) f1s[t] = to_f32(fmap1[o1 + t]);
  for (int t = threadIdx.x; t < 2 * PP; t += blockDim.x) cxy[t] = coords[(size_t)be * 2 * PP + t];
  __syncthreads();
  const float* g = grad + (size_t)be * TT * TT * PP;
  for (int t = threadIdx.x; t < PP * D * D; t += blockDim.x) {
    const int b = t % D, a = (t / D) % D, p = t / (D * D);
    const float x = cxy[p], y = cxy[PP + p];
    const float dx = x - floorf(x), dy = y - floorf(y);
    float s = 0.f;  // gather form of the four zero-padded index_put_ tensors g1..g4
    if (a < TT && b < TT) s += (1.f - dx) * (1.f - dy) * g[(b * TT + a) * PP + p];
    if (a < TT && b > 0) s += dx * (1.f - dy) * g[((b - 1) * TT + a) * PP + p];
    if (a > 0 && b < TT) s += (1.f - dx) * dy * g[(b * TT + a - 1) * PP + p];
    if (a > 0 && b > 0) s += dx * dy * g[((b - 1) * TT + a - 1) * PP + p];
    cg[t] = s;
  }
  __syncthreads();
  const size_t plane = (size_t)H2 * W2;
  // fmap2_grad: one atomic per (channel, tap); fmap1_grad: taps reduced per (channel, pixel) first
  for (int t = threadIdx.x; t < C * PP; t += blockDim.x) {
    const int p = t % PP, c = t / PP;
    const int fy = (int)floorf(cxy[PP + p]), fx = (int)floorf(cxy[p]);
    const float f1 = f1s[c * PP + p];
    float s1 = 0.f;
    for (int a = 0; a < D; a++) {
      const int i1 = fy + a - R;
      if (i1 < 0 || i1 >= H2) continue;
      for (int b = 0; b < D; b++) {
        const int j1 = fx + b - R;
        if (j1 < 0 || j1 >= W2) continue;
        const float gv = cg[(p * D + a) * D + b];
        const size_t a2 = o2 + (size_t)c * plane + (size_t)i1 * W2 + j1;
        s1 = fmaf(gv, to_f32(fmap2[a2]), s1);
        atomicAdd(fmap2_grad + a2, gv * f1);
      }
    }
    atomicAdd(fmap1_grad + o1 + t, s1);
  }
}

int corr_backward_generic(const void* fmap1, const void* fmap2, const float* coords, const int64_t* ii,
                          const int64_t* jj, const float* grad, float* g1, float* g2, int dtype, int B, int E,
                          int C, int PH, int PW, int n1, int n2, int H2, int W2, int R, int* status,
                          cudaStream_t st) {
  const int PP = PH * PW, D = 2 * R + 2;
  const size_t smem = sizeof(float) * ((size_t)C * PP + (size_t)PP * D * D + 2 * PP);
  if (smem > 200 * 1024) return WV_EUNSUPPORTED;
  if (dtype == WV_F16) {
    if (smem > 48 * 1024)
      WV_CUDA_TRY(cudaFuncSetAttribute(corr_bwd_generic_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    corr_bwd_generic_kernel<__half><<<B * E, 128, smem, st>>>((const __half*)fmap1, (const __half*)fmap2, coords, ii, jj,
                                                             grad, g1, g2, E, C, PH, PW, n1, n2, H2, W2, R, status);
  } else {
    if (smem > 48 * 1024)
      WV_CUDA_TRY(cudaFuncSetAttribute(corr_bwd_generic_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    corr_bwd_generic_kernel<float><<<B * E, 128, smem, st>>>((const float*)fmap1, (const float*)fmap2, coords, ii, jj,
                                                            grad, g1, g2, E, C, PH, PW, n1, n2, H2, W2, R, status);
  }
  WV_LAUNCH_CHECK();
  return WV_OK;
}

// ---------------------------------------------------------------------------------------------------
// patchify: correlation_kernel.cu:16-47 (forward), :49-80 (backward)
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void patchify_fwd_kernel(const T* __restrict__ net, const float* __restrict__ coords,
                                    T* __restrict__ patches, int B, int M, int C, int H, int W, int R) {
  const int D = 2 * R + 2;
  const size_t total = (size_t)B * M * C * D * D;
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    size_t r = t;
    const int b = r % D; r /= D;
    const int a = r % D; r /= D;
    const int k = r % C; r /= C;
    const int m = r % M; const int n = (int)(r / M);
    const float x = coords[((size_t)n * M + m) * 2], y = coords[((size_t)n * M + m) * 2 + 1];
    const int i = (int)floorf(y) + (a - R), j = (int)floorf(x) + (b - R);
    T v = from_f32<T>(0.f);
    if (i >= 0 && i < H && j >= 0 && j < W) v = net[(((size_t)n * C + k) * H + i) * W + j];
    patches[t] = v;
  }
}

template <typename T>
__global__ void patchify_bwd_kernel(const T* __restrict__ gradient, const float* __restrict__ coords,
                                    float* __restrict__ net_grad, int B, int M, int C, int H, int W, int R) {
  const int D = 2 * R + 2;
  const size_t total = (size_t)B * M * C * D * D;
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    size_t r = t;
    const int b = r % D; r /= D;
    const int a = r % D; r /= D;
    const int k = r % C; r /= C;
    const int m = r % M; const int n = (int)(r / M);
    const float x = coords[((size_t)n * M + m) * 2], y = coords[((size_t)n * M + m) * 2 + 1];
    const int i = (int)floorf(y) + (a - R), j = (int)floorf(x) + (b - R);
    if (i >= 0 && i < H && j >= 0 && j < W)
      atomicAdd(net_grad + (((size_t)n * C + k) * H + i) * W + j, to_f32(gradient[t]));
  }
}

static int grid_for(size_t total, int threads) {
  size_t g = (total + threads - 1) / threads;
  const size_t cap = (size_t)sm_count() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

int patchify_forward(const void* net, const float* coords, void* patches, int dtype, int B, int M, int C, int H,
                     int W, int R, cudaStream_t st) {
  const int D = 2 * R + 2;
  const size_t total = (size_t)B * M * C * D * D;
  if (total == 0) return WV_OK;
  if (dtype == WV_F16)
    patchify_fwd_kernel<__half><<<grid_for(total, 256), 256, 0, st>>>((const __half*)net, coords, (__half*)patches, B, M, C, H, W, R);
  else
    patchify_fwd_kernel<float><<<grid_for(total, 256), 256, 0, st>>>((const float*)net, coords, (float*)patches, B, M, C, H, W, R);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

int patchify_backward(const void* gradient, const float* coords, float* net_grad, int dtype, int B, int M, int C,
                      int H, int W, int R, cudaStream_t st) {
  const int D = 2 * R + 2;
  const size_t total = (size_t)B * M * C * D * D;
  if (total == 0) return WV_OK;
  if (dtype == WV_F16)
    patchify_bwd_kernel<__half><<<grid_for(total, 256), 256, 0, st>>>((const __half*)gradient, coords, net_grad, B, M, C, H, W, R);
  else
    patchify_bwd_kernel<float><<<grid_for(total, 256), 256, 0, st>>>((const float*)gradient, coords, net_grad, B, M, C, H, W, R);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

}  // namespace wv
