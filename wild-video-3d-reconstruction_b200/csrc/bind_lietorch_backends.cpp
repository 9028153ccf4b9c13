// torch extension `lietorch_backends`: the reference's module name and m.def table (dpvo/lietorch/src/lietorch.cpp:286-316,
// imported by dpvo/lietorch/group_ops.py:1) over wv_lie_forward / wv_lie_backward of libwvb200.so.  Inputs arrive flattened to
// [batch, dim] and contiguous (dpvo/lietorch/broadcasting.py:9-31); CUDA tensors run the sm_100a kernels on the current stream,
// CPU tensors run the same algebra on the host like the reference's lietorch_cpu.cpp.  float and double.
#include "bind_common.h"

namespace {

using torch::Tensor;

struct Dims { int K, N; };
Dims dims_of(int group) {
  switch (group) {
    case WV_LIE_SO3: return {3, 4};
    case WV_LIE_RXSO3: return {4, 5};
    case WV_LIE_SE3: return {6, 7};
    case WV_LIE_SIM3: return {7, 8};
  }
  TORCH_CHECK(false, "lietorch_backends: unknown group id ", group);
  return {0, 0};
}

int lie_dtype(const Tensor& t, const char* name) {
  if (t.scalar_type() == torch::kFloat) return WV_F32;
  if (t.scalar_type() == torch::kDouble) return WV_F64;
  TORCH_CHECK(false, name, ": expected a float or double tensor, got ", t.scalar_type());
  return -1;
}

void check_arg(const Tensor& t, const char* name, int64_t cols, const Tensor& like) {
  TORCH_CHECK(t.is_contiguous(), name, " must be contiguous");
  TORCH_CHECK(t.dim() == 2 && t.size(1) == cols, name, ": expected [batch, ", cols, "], got ", t.sizes());
  TORCH_CHECK(t.device() == like.device() && t.scalar_type() == like.scalar_type(), name, ": device / dtype differ between the arguments");
  TORCH_CHECK(t.size(0) == like.size(0), name, ": batch sizes differ (", t.size(0), " vs ", like.size(0), ")");
}

struct Launch {
  bool host;
  wv_stream_t stream;
  c10::cuda::OptionalCUDAGuard guard;
  explicit Launch(const Tensor& t) : host(!t.is_cuda()), stream(nullptr) {
    TORCH_CHECK(t.is_cuda() || t.is_cpu(), "lietorch_backends: CPU or CUDA tensors only");
    if (!host) {
      guard.set_device(t.device());
      stream = wvbind::current_stream();
    }
  }
};

Tensor forward(int op, int group, const Tensor& X, int64_t xcols, const Tensor* Y, int64_t ycols, std::vector<int64_t> out_shape,
               const char* what) {
  check_arg(X, "X", xcols, X);
  if (Y) check_arg(*Y, "Y", ycols, X);
  Launch L(X);
  Tensor out = torch::empty(out_shape, X.options());
  wvbind::check_rc(wv_lie_forward(op, group, lie_dtype(X, what), X.data_ptr(), xcols, Y ? Y->data_ptr() : nullptr, ycols, out.data_ptr(),
                                  X.size(0), L.host, L.stream), what);
  return out;
}

std::vector<Tensor> backward(int op, int group, const Tensor& grad, int64_t gcols, const Tensor& X, int64_t xcols, const Tensor* Y,
                             int64_t ycols, const char* what) {
  check_arg(X, "X", xcols, X);
  TORCH_CHECK(grad.is_contiguous(), "grad must be contiguous");
  TORCH_CHECK(grad.numel() == X.size(0) * gcols && grad.device() == X.device() && grad.scalar_type() == X.scalar_type(),
              what, ": grad does not match the forward output");
  if (Y) check_arg(*Y, "Y", ycols, X);
  Launch L(X);
  Tensor dX = torch::empty_like(X);
  Tensor dY = Y ? torch::empty_like(*Y) : Tensor();
  wvbind::check_rc(wv_lie_backward(op, group, lie_dtype(X, what), grad.data_ptr(), X.data_ptr(), xcols, Y ? Y->data_ptr() : nullptr, ycols,
                                   dX.data_ptr(), Y ? dY.data_ptr() : nullptr, X.size(0), L.host, L.stream), what);
  if (Y) return {dX, dY};
  return {dX};
}

// unary
Tensor expm(int g, Tensor a) { auto d = dims_of(g); return forward(WV_LIE_EXP, g, a, d.K, nullptr, 0, {a.size(0), d.N}, "expm"); }
std::vector<Tensor> expm_backward(int g, Tensor grad, Tensor a) { auto d = dims_of(g); return backward(WV_LIE_EXP, g, grad, d.N, a, d.K, nullptr, 0, "expm_backward"); }
Tensor logm(int g, Tensor X) { auto d = dims_of(g); return forward(WV_LIE_LOG, g, X, d.N, nullptr, 0, {X.size(0), d.K}, "logm"); }
std::vector<Tensor> logm_backward(int g, Tensor grad, Tensor X) { auto d = dims_of(g); return backward(WV_LIE_LOG, g, grad, d.K, X, d.N, nullptr, 0, "logm_backward"); }
Tensor inv(int g, Tensor X) { auto d = dims_of(g); return forward(WV_LIE_INV, g, X, d.N, nullptr, 0, {X.size(0), d.N}, "inv"); }
std::vector<Tensor> inv_backward(int g, Tensor grad, Tensor X) { auto d = dims_of(g); return backward(WV_LIE_INV, g, grad, d.N, X, d.N, nullptr, 0, "inv_backward"); }
// binary
Tensor mul(int g, Tensor X, Tensor Y) { auto d = dims_of(g); return forward(WV_LIE_MUL, g, X, d.N, &Y, d.N, {X.size(0), d.N}, "mul"); }
std::vector<Tensor> mul_backward(int g, Tensor grad, Tensor X, Tensor Y) { auto d = dims_of(g); return backward(WV_LIE_MUL, g, grad, d.N, X, d.N, &Y, d.N, "mul_backward"); }
Tensor adj(int g, Tensor X, Tensor a) { auto d = dims_of(g); return forward(WV_LIE_ADJ, g, X, d.N, &a, d.K, {X.size(0), d.K}, "adj"); }
std::vector<Tensor> adj_backward(int g, Tensor grad, Tensor X, Tensor a) { auto d = dims_of(g); return backward(WV_LIE_ADJ, g, grad, d.K, X, d.N, &a, d.K, "adj_backward"); }
Tensor adjT(int g, Tensor X, Tensor a) { auto d = dims_of(g); return forward(WV_LIE_ADJT, g, X, d.N, &a, d.K, {X.size(0), d.K}, "adjT"); }
std::vector<Tensor> adjT_backward(int g, Tensor grad, Tensor X, Tensor a) { auto d = dims_of(g); return backward(WV_LIE_ADJT, g, grad, d.K, X, d.N, &a, d.K, "adjT_backward"); }
Tensor act(int g, Tensor X, Tensor p) { auto d = dims_of(g); return forward(WV_LIE_ACT, g, X, d.N, &p, 3, {X.size(0), 3}, "act"); }
std::vector<Tensor> act_backward(int g, Tensor grad, Tensor X, Tensor p) { auto d = dims_of(g); return backward(WV_LIE_ACT, g, grad, 3, X, d.N, &p, 3, "act_backward"); }
Tensor act4(int g, Tensor X, Tensor p) { auto d = dims_of(g); return forward(WV_LIE_ACT4, g, X, d.N, &p, 4, {X.size(0), 4}, "act4"); }
std::vector<Tensor> act4_backward(int g, Tensor grad, Tensor X, Tensor p) { auto d = dims_of(g); return backward(WV_LIE_ACT4, g, grad, 4, X, d.N, &p, 4, "act4_backward"); }
// no gradient
Tensor as_matrix(int g, Tensor X) { auto d = dims_of(g); return forward(WV_LIE_MATRIX, g, X, d.N, nullptr, 0, {X.size(0), 4, 4}, "as_matrix"); }
Tensor projector(int g, Tensor X) { auto d = dims_of(g); return forward(WV_LIE_PROJECTOR, g, X, d.N, nullptr, 0, {X.size(0), d.N, d.N}, "projector"); }
Tensor Jinv(int g, Tensor X, Tensor a) { auto d = dims_of(g); return forward(WV_LIE_JINV, g, X, d.N, &a, d.K, {X.size(0), d.K}, "Jinv"); }

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("expm", &expm, "exp map forward");
  m.def("expm_backward", &expm_backward, "exp map backward");
  m.def("logm", &logm, "log map forward");
  m.def("logm_backward", &logm_backward, "log map backward");
  m.def("inv", &inv, "inverse operator");
  m.def("inv_backward", &inv_backward, "inverse operator backward");
  m.def("mul", &mul, "group operator");
  m.def("mul_backward", &mul_backward, "group operator backward");
  m.def("adj", &adj, "adjoint operator");
  m.def("adj_backward", &adj_backward, "adjoint operator backward");
  m.def("adjT", &adjT, "transposed adjoint operator");
  m.def("adjT_backward", &adjT_backward, "transposed adjoint operator backward");
  m.def("act", &act, "action on point");
  m.def("act_backward", &act_backward, "action on point backward");
  m.def("act4", &act4, "action on homogeneous point");
  m.def("act4_backward", &act4_backward, "action on homogeneous point backward");
  m.def("as_matrix", &as_matrix, "convert to matrix");
  m.def("projector", &projector, "orthogonal projection matrix");
  m.def("Jinv", &Jinv, "left inverse jacobian operator");
}
