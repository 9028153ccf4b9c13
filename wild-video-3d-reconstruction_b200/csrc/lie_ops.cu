// wv_lie_forward / wv_lie_backward: the group operations behind `lietorch_backends` (include/wvb200.h), one group element per
// thread.  Replaces the reference's 19 Eigen-based kernels (dpvo/lietorch/src/lietorch_gpu.cu:20-294) and their at::parallel_for
// twins (lietorch_cpu.cpp:15-352) with one templated body over the plain-array algebra of lie.cuh, compiled for device and host.
//
// Per-argument batch strides (in scalars) let a caller pass a broadcast operand once (stride 0) instead of materialising it with
// repeat() as dpvo/lietorch/broadcasting.py:26-29 does; the drop-in binding passes dense strides because that Python layer has
// already repeated.  These are latency-sized launches (E or 9 E elements of <= 8 scalars); the grid is capped at 8 CTAs per SM.
#include <stdint.h>

#include "lie.cuh"
#include "wv_common.cuh"

namespace wv {
namespace {

using namespace lie;

template <typename T, int N>
LIE_HD Vec<T, N> load(const T* p) {
  Vec<T, N> v;
#pragma unroll
  for (int i = 0; i < N; ++i) v[i] = p[i];
  return v;
}
template <typename T, int N>
LIE_HD void store(T* p, const Vec<T, N>& v) {
#pragma unroll
  for (int i = 0; i < N; ++i) p[i] = v[i];
}
template <typename T, int R, int C>
LIE_HD void store(T* p, const Mat<T, R, C>& m) {
#pragma unroll
  for (int i = 0; i < R * C; ++i) p[i] = m.a[i];
}

// output row length of a forward op / of the second operand
template <class G> LIE_HD constexpr int out_len(int op) {
  return op == WV_LIE_EXP || op == WV_LIE_INV || op == WV_LIE_MUL ? G::N
         : op == WV_LIE_LOG || op == WV_LIE_ADJ || op == WV_LIE_ADJT || op == WV_LIE_JINV ? G::K
         : op == WV_LIE_ACT ? 3 : op == WV_LIE_ACT4 ? 4 : op == WV_LIE_MATRIX ? 16 : G::N * G::N;
}

template <class G, typename T, int OP>
LIE_HD void forward_one(const T* x, const T* y, T* o) {
  if constexpr (OP == WV_LIE_EXP) {
    G::Exp(load<T, G::K>(x)).store(o);
  } else if constexpr (OP == WV_LIE_LOG) {
    store(o, G(x).Log());
  } else if constexpr (OP == WV_LIE_INV) {
    G(x).inv().store(o);
  } else if constexpr (OP == WV_LIE_MUL) {
    (G(x) * G(y)).store(o);
  } else if constexpr (OP == WV_LIE_ADJ) {
    store(o, G(x).Adj() * load<T, G::K>(y));
  } else if constexpr (OP == WV_LIE_ADJT) {
    store(o, transpose(G(x).Adj()) * load<T, G::K>(y));
  } else if constexpr (OP == WV_LIE_ACT) {
    store(o, G(x).act(load<T, 3>(y)));
  } else if constexpr (OP == WV_LIE_ACT4) {
    store(o, G(x).act4(load<T, 4>(y)));
  } else if constexpr (OP == WV_LIE_MATRIX) {
    store(o, G(x).matrix4());
  } else if constexpr (OP == WV_LIE_PROJECTOR) {
    store(o, G(x).projector());
  } else if constexpr (OP == WV_LIE_JINV) {
    store(o, G::left_jacobian_inverse(G(x).Log()) * load<T, G::K>(y));
  }
}

// gradients are row vectors over the first K entries of an N-wide row, exactly as the reference maps them
// (lietorch_gpu.cu:33-43: Grad = Matrix<1, K> read from grad + i * N); entries K..N-1 of dX stay zero (written here).
template <class G, typename T>
LIE_HD void store_group_grad(T* dX, const Vec<T, G::K>& g) {
#pragma unroll
  for (int i = 0; i < G::K; ++i) dX[i] = g[i];
#pragma unroll
  for (int i = G::K; i < G::N; ++i) dX[i] = T(0);
}

template <class G, typename T, int OP>
LIE_HD void backward_one(const T* grad, const T* x, const T* y, T* dx, T* dy) {
  constexpr int K = G::K;
  if constexpr (OP == WV_LIE_EXP) {
    store(dx, rowmul(load<T, K>(grad), G::left_jacobian(load<T, K>(x))));
  } else if constexpr (OP == WV_LIE_LOG) {
    store_group_grad<G, T>(dx, rowmul(load<T, K>(grad), G::left_jacobian_inverse(G(x).Log())));
  } else if constexpr (OP == WV_LIE_INV) {
    store_group_grad<G, T>(dx, -rowmul(load<T, K>(grad), G(x).inv().Adj()));
  } else if constexpr (OP == WV_LIE_MUL) {
    const Vec<T, K> dz = load<T, K>(grad);
    store_group_grad<G, T>(dx, dz);
    store_group_grad<G, T>(dy, rowmul(dz, G(x).Adj()));
  } else if constexpr (OP == WV_LIE_ADJ) {
    const G X(x);
    const Vec<T, K> db = load<T, K>(grad);
    const typename G::Adjoint A = X.Adj();
    const Vec<T, K> b = A * load<T, K>(y);
    store(dy, rowmul(db, A));
    store_group_grad<G, T>(dx, -rowmul(db, G::adj(b)));
  } else if constexpr (OP == WV_LIE_ADJT) {
    const G X(x);
    const Vec<T, K> c = X.Adj() * load<T, K>(grad);
    store(dy, c);
    store_group_grad<G, T>(dx, -rowmul(load<T, K>(y), G::adj(c)));
  } else if constexpr (OP == WV_LIE_ACT) {
    const G X(x);
    const Vec<T, 3> dq = load<T, 3>(grad);
    const Mat<T, 4, 4> M = X.matrix4();
    Vec<T, 3> dp;
#pragma unroll
    for (int c = 0; c < 3; ++c) dp[c] = dq[0] * M(0, c) + dq[1] * M(1, c) + dq[2] * M(2, c);
    store(dy, dp);
    store_group_grad<G, T>(dx, rowmul(dq, G::act_jacobian(X.act(load<T, 3>(y)))));
  } else if constexpr (OP == WV_LIE_ACT4) {
    const G X(x);
    const Vec<T, 4> dq = load<T, 4>(grad);
    store(dy, rowmul(dq, X.matrix4()));
    store_group_grad<G, T>(dx, rowmul(dq, G::act4_jacobian(X.act4(load<T, 4>(y)))));
  }
}

template <class G, typename T, int OP>
__global__ void __launch_bounds__(256) lie_forward_kernel(const T* __restrict__ X, int64_t xs, const T* __restrict__ Y, int64_t ys,
                                                          T* __restrict__ out, int64_t n) {
  constexpr int L = out_len<G>(OP);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    forward_one<G, T, OP>(X + i * xs, Y ? Y + i * ys : nullptr, out + i * L);
}

template <class G, typename T, int OP>
__global__ void __launch_bounds__(256) lie_backward_kernel(const T* __restrict__ grad, const T* __restrict__ X, int64_t xs,
                                                           const T* __restrict__ Y, int64_t ys, T* __restrict__ dX, T* __restrict__ dY,
                                                           int64_t n) {
  constexpr int LG = out_len<G>(OP);                                     // grad row = forward output row
  constexpr int LX = OP == WV_LIE_EXP ? G::K : G::N;                     // dX row
  constexpr int LY = OP == WV_LIE_MUL ? G::N : OP == WV_LIE_ACT ? 3 : OP == WV_LIE_ACT4 ? 4 : G::K;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    backward_one<G, T, OP>(grad + i * LG, X + i * xs, Y ? Y + i * ys : nullptr, dX + i * LX, dY ? dY + i * LY : nullptr);
}

template <class G, typename T, int OP>
int run_forward(const void* X, int64_t xs, const void* Y, int64_t ys, void* out, int64_t n, bool host, cudaStream_t st) {
  constexpr int L = out_len<G>(OP);
  const T* x = static_cast<const T*>(X);
  const T* y = static_cast<const T*>(Y);
  T* o = static_cast<T*>(out);
  if (host) {
    for (int64_t i = 0; i < n; ++i) forward_one<G, T, OP>(x + i * xs, y ? y + i * ys : nullptr, o + i * L);
    return WV_OK;
  }
  const int64_t want = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  lie_forward_kernel<G, T, OP><<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(x, xs, y, ys, o, n);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

template <class G, typename T, int OP>
int run_backward(const void* grad, const void* X, int64_t xs, const void* Y, int64_t ys, void* dX, void* dY, int64_t n, bool host,
                 cudaStream_t st) {
  constexpr int LG = out_len<G>(OP);
  constexpr int LX = OP == WV_LIE_EXP ? G::K : G::N;
  constexpr int LY = OP == WV_LIE_MUL ? G::N : OP == WV_LIE_ACT ? 3 : OP == WV_LIE_ACT4 ? 4 : G::K;
  const T* g = static_cast<const T*>(grad);
  const T* x = static_cast<const T*>(X);
  const T* y = static_cast<const T*>(Y);
  T* dx = static_cast<T*>(dX);
  T* dy = static_cast<T*>(dY);
  if (host) {
    for (int64_t i = 0; i < n; ++i)
      backward_one<G, T, OP>(g + i * LG, x + i * xs, y ? y + i * ys : nullptr, dx + i * LX, dy ? dy + i * LY : nullptr);
    return WV_OK;
  }
  const int64_t want = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  lie_backward_kernel<G, T, OP><<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(g, x, xs, y, ys, dx, dy, n);
  WV_LAUNCH_CHECK();
  return WV_OK;
}

template <class G, typename T>
int dispatch_forward(int op, const void* X, int64_t xs, const void* Y, int64_t ys, void* out, int64_t n, bool host, cudaStream_t st) {
  switch (op) {
#define WV_CASE(OP) case OP: return run_forward<G, T, OP>(X, xs, Y, ys, out, n, host, st)
    WV_CASE(WV_LIE_EXP); WV_CASE(WV_LIE_LOG); WV_CASE(WV_LIE_INV); WV_CASE(WV_LIE_MUL); WV_CASE(WV_LIE_ADJ); WV_CASE(WV_LIE_ADJT);
    WV_CASE(WV_LIE_ACT); WV_CASE(WV_LIE_ACT4); WV_CASE(WV_LIE_MATRIX); WV_CASE(WV_LIE_PROJECTOR); WV_CASE(WV_LIE_JINV);
#undef WV_CASE
  }
  return WV_EINVAL;
}
template <class G, typename T>
int dispatch_backward(int op, const void* grad, const void* X, int64_t xs, const void* Y, int64_t ys, void* dX, void* dY, int64_t n,
                      bool host, cudaStream_t st) {
  switch (op) {
#define WV_CASE(OP) case OP: return run_backward<G, T, OP>(grad, X, xs, Y, ys, dX, dY, n, host, st)
    WV_CASE(WV_LIE_EXP); WV_CASE(WV_LIE_LOG); WV_CASE(WV_LIE_INV); WV_CASE(WV_LIE_MUL); WV_CASE(WV_LIE_ADJ); WV_CASE(WV_LIE_ADJT);
    WV_CASE(WV_LIE_ACT); WV_CASE(WV_LIE_ACT4);
#undef WV_CASE
  }
  return WV_EINVAL;  // as_matrix / projector / Jinv have no backward in the reference either (group_ops.py:73-79)
}

template <typename T>
int by_group_forward(int group, int op, const void* X, int64_t xs, const void* Y, int64_t ys, void* out, int64_t n, bool host,
                     cudaStream_t st) {
  switch (group) {
    case WV_LIE_SO3: return dispatch_forward<SO3<T>, T>(op, X, xs, Y, ys, out, n, host, st);
    case WV_LIE_RXSO3: return dispatch_forward<RxSO3<T>, T>(op, X, xs, Y, ys, out, n, host, st);
    case WV_LIE_SE3: return dispatch_forward<SE3<T>, T>(op, X, xs, Y, ys, out, n, host, st);
    case WV_LIE_SIM3: return dispatch_forward<Sim3<T>, T>(op, X, xs, Y, ys, out, n, host, st);
  }
  return WV_EINVAL;
}
template <typename T>
int by_group_backward(int group, int op, const void* grad, const void* X, int64_t xs, const void* Y, int64_t ys, void* dX, void* dY,
                      int64_t n, bool host, cudaStream_t st) {
  switch (group) {
    case WV_LIE_SO3: return dispatch_backward<SO3<T>, T>(op, grad, X, xs, Y, ys, dX, dY, n, host, st);
    case WV_LIE_RXSO3: return dispatch_backward<RxSO3<T>, T>(op, grad, X, xs, Y, ys, dX, dY, n, host, st);
    case WV_LIE_SE3: return dispatch_backward<SE3<T>, T>(op, grad, X, xs, Y, ys, dX, dY, n, host, st);
    case WV_LIE_SIM3: return dispatch_backward<Sim3<T>, T>(op, grad, X, xs, Y, ys, dX, dY, n, host, st);
  }
  return WV_EINVAL;
}

bool binary_op(int op) { return op == WV_LIE_MUL || op == WV_LIE_ADJ || op == WV_LIE_ADJT || op == WV_LIE_ACT || op == WV_LIE_ACT4 || op == WV_LIE_JINV; }

}  // namespace
}  // namespace wv

extern "C" {

int wv_lie_forward(int op, int group, int dtype, const void* X, int64_t x_stride, const void* Y, int64_t y_stride, void* out, int64_t n,
                   int on_host, wv_stream_t stream) {
  if (n < 0 || (n > 0 && (!X || !out)) || (wv::binary_op(op) && n > 0 && !Y)) return WV_EINVAL;
  if (n == 0) return WV_OK;
  if (!wv::binary_op(op)) Y = nullptr;
  if (dtype == WV_F32) return wv::by_group_forward<float>(group, op, X, x_stride, Y, y_stride, out, n, on_host != 0, wv::as_stream(stream));
  if (dtype == WV_F64) return wv::by_group_forward<double>(group, op, X, x_stride, Y, y_stride, out, n, on_host != 0, wv::as_stream(stream));
  return WV_EINVAL;
}

int wv_lie_backward(int op, int group, int dtype, const void* grad, const void* X, int64_t x_stride, const void* Y, int64_t y_stride,
                    void* dX, void* dY, int64_t n, int on_host, wv_stream_t stream) {
  const bool bin = wv::binary_op(op);
  if (n < 0 || (n > 0 && (!grad || !X || !dX)) || (bin && n > 0 && (!Y || !dY))) return WV_EINVAL;
  if (n == 0) return WV_OK;
  if (!bin) { Y = nullptr; dY = nullptr; }
  if (dtype == WV_F32) return wv::by_group_backward<float>(group, op, grad, X, x_stride, Y, y_stride, dX, dY, n, on_host != 0, wv::as_stream(stream));
  if (dtype == WV_F64) return wv::by_group_backward<double>(group, op, grad, X, x_stride, Y, y_stride, dX, dY, n, on_host != 0, wv::as_stream(stream));
  return WV_EINVAL;
}

}  // extern "C"
