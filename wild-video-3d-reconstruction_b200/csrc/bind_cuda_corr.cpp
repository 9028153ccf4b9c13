// torch extension `cuda_corr`: same module name and m.def table as the reference's
// dpvo/altcorr/correlation.cpp:57-63, so dpvo/altcorr/correlation.py:2 imports it unchanged.
#include "bind_common.h"

#include <cmath>

using namespace wvbind;

namespace {

// double maps are accepted like the reference's AT_DISPATCH_FLOATING_TYPES_AND_HALF does; computed in fp32
torch::Tensor as_map(const torch::Tensor& t) {
  return (t.scalar_type() == torch::kDouble ? t.to(torch::kFloat) : t).contiguous();
}

std::vector<torch::Tensor> corr_forward(torch::Tensor fmap1, torch::Tensor fmap2, torch::Tensor coords, torch::Tensor ii,
                                        torch::Tensor jj, int radius) {
  check_cuda(fmap1, "fmap1"); check_cuda(fmap2, "fmap2"); check_cuda(coords, "coords"); check_cuda(ii, "ii"); check_cuda(jj, "jj");
  TORCH_CHECK(fmap1.dim() == 5 && fmap2.dim() == 5 && coords.dim() == 5, "corr: fmap1/fmap2/coords must be 5-d");
  TORCH_CHECK(fmap1.scalar_type() == fmap2.scalar_type(), "corr: fmap1 and fmap2 must have the same dtype");
  TORCH_CHECK(coords.scalar_type() == torch::kFloat, "corr: coords must be float32");
  TORCH_CHECK(ii.scalar_type() == torch::kLong && jj.scalar_type() == torch::kLong, "corr: ii/jj must be int64");
  c10::cuda::CUDAGuard guard(fmap1.device());
  const auto in_dtype = fmap1.scalar_type();
  torch::Tensor f1 = as_map(fmap1), f2 = as_map(fmap2), co = coords.contiguous(), iic = ii.contiguous(), jjc = jj.contiguous();
  const int dt = wv_dtype_of(f1, "corr");
  const int B = co.size(0), E = co.size(1), PH = co.size(3), PW = co.size(4);
  const int n1 = f1.size(1), C = f1.size(2), n2 = f2.size(1), H2 = f2.size(3), W2 = f2.size(4);
  TORCH_CHECK(co.size(2) == 2 && f1.size(3) == PH && f1.size(4) == PW && f2.size(2) == C && f1.size(0) == B && f2.size(0) == B,
              "corr: inconsistent shapes");
  TORCH_CHECK(iic.numel() == E && jjc.numel() == E, "corr: ii/jj must have one entry per edge");
  const int T = 2 * radius + 1;
  torch::Tensor out = torch::empty({B, E, T, T, PH, PW}, f1.options());
  const size_t wsb = wv_corr_workspace_bytes(dt, B, E, C, PH, PW, n1, n2, H2, W2, radius);
  torch::Tensor ws;
  if (wsb) ws = torch::empty({(int64_t)wsb}, torch::dtype(torch::kUInt8).device(f1.device()));
  check_rc(wv_corr_forward(f1.data_ptr(), f2.data_ptr(), co.data_ptr<float>(), iic.data_ptr<int64_t>(), jjc.data_ptr<int64_t>(),
                           out.data_ptr(), dt, B, E, C, PH, PW, n1, n2, H2, W2, radius, wsb ? ws.data_ptr() : nullptr, wsb,
                           status_word(f1.device()).data_ptr<int>(), current_stream()),
           "cuda_corr.forward");
  if (env_flag("WVB200_CHECK_INDICES")) raise_on_status(f1.device(), "cuda_corr.forward");
  if (in_dtype == torch::kDouble) out = out.to(torch::kDouble);
  return {out};
}

// DPVO.corr (dpvo/dpvo.py:325-333) in one call: returns [B, E, T*T*PH*PW*2] == torch.stack([corr1, corr2], -1).view(B, E, -1).
// div0 / div1 are the reference's coordinate divisors (1 and 4).  Raises for configurations the two-level kernel does not cover;
// the Python mirror (altcorr.corr_stacked) composes those from forward().
std::vector<torch::Tensor> corr_forward_stacked(torch::Tensor fmap1, torch::Tensor fmap2_l0, torch::Tensor fmap2_l1,
                                                torch::Tensor coords, torch::Tensor ii, torch::Tensor jj, int radius, double div0,
                                                double div1) {
  check_cuda(fmap1, "fmap1"); check_cuda(fmap2_l0, "fmap2_l0"); check_cuda(fmap2_l1, "fmap2_l1"); check_cuda(coords, "coords");
  check_cuda(ii, "ii"); check_cuda(jj, "jj");
  TORCH_CHECK(fmap1.dim() == 5 && fmap2_l0.dim() == 5 && fmap2_l1.dim() == 5 && coords.dim() == 5, "corr: fmap1/fmap2/coords must be 5-d");
  TORCH_CHECK(fmap1.scalar_type() == torch::kHalf && fmap2_l0.scalar_type() == torch::kHalf && fmap2_l1.scalar_type() == torch::kHalf,
              "corr_stacked: half maps only");
  TORCH_CHECK(coords.scalar_type() == torch::kFloat, "corr: coords must be float32");
  TORCH_CHECK(ii.scalar_type() == torch::kLong && jj.scalar_type() == torch::kLong, "corr: ii/jj must be int64");
  int e0 = 0, e1 = 0;
  TORCH_CHECK(std::frexp(div0, &e0) == 0.5 && std::frexp(div1, &e1) == 0.5, "corr_stacked: divisors must be powers of two");
  c10::cuda::CUDAGuard guard(fmap1.device());
  torch::Tensor f1 = fmap1.contiguous(), f20 = fmap2_l0.contiguous(), f21 = fmap2_l1.contiguous(), co = coords.contiguous(),
                iic = ii.contiguous(), jjc = jj.contiguous();
  const int B = co.size(0), E = co.size(1), PH = co.size(3), PW = co.size(4);
  const int n1 = f1.size(1), C = f1.size(2), n2 = f20.size(1), H0 = f20.size(3), W0 = f20.size(4), H1 = f21.size(3), W1 = f21.size(4);
  TORCH_CHECK(co.size(2) == 2 && f1.size(3) == PH && f1.size(4) == PW && f20.size(2) == C && f21.size(2) == C && f21.size(1) == n2 &&
                  f1.size(0) == B && f20.size(0) == B && f21.size(0) == B,
              "corr_stacked: inconsistent shapes");
  TORCH_CHECK(iic.numel() == E && jjc.numel() == E, "corr: ii/jj must have one entry per edge");
  const int T = 2 * radius + 1;
  const size_t wsb = wv_corr_stacked_workspace_bytes(WV_F16, B, E, C, PH, PW, n1, n2, H0, W0, H1, W1, radius);
  torch::Tensor out = torch::empty({B, E, (int64_t)T * T * PH * PW * 2}, f1.options());
  if (E == 0) return {out};
  TORCH_CHECK(wsb > 0, "corr_stacked: configuration not covered by the two-level kernel");
  torch::Tensor ws = torch::empty({(int64_t)wsb}, torch::dtype(torch::kUInt8).device(f1.device()));
  check_rc(wv_corr_forward_stacked(f1.data_ptr(), f20.data_ptr(), f21.data_ptr(), co.data_ptr<float>(), (float)(1.0 / div0),
                                   (float)(1.0 / div1), iic.data_ptr<int64_t>(), jjc.data_ptr<int64_t>(), out.data_ptr(), WV_F16, B, E,
                                   C, PH, PW, n1, n2, H0, W0, H1, W1, radius, ws.data_ptr(), wsb,
                                   status_word(f1.device()).data_ptr<int>(), current_stream()),
           "cuda_corr.forward_stacked");
  if (env_flag("WVB200_CHECK_INDICES")) raise_on_status(f1.device(), "cuda_corr.forward_stacked");
  return {out};
}

std::vector<torch::Tensor> corr_backward(torch::Tensor fmap1, torch::Tensor fmap2, torch::Tensor coords, torch::Tensor ii,
                                         torch::Tensor jj, torch::Tensor corr_grad, int radius) {
  check_cuda(fmap1, "fmap1"); check_cuda(fmap2, "fmap2"); check_cuda(coords, "coords"); check_cuda(corr_grad, "corr_grad");
  c10::cuda::CUDAGuard guard(fmap1.device());
  const auto in_dtype = fmap1.scalar_type();
  torch::Tensor f1 = as_map(fmap1), f2 = as_map(fmap2), co = coords.contiguous(), iic = ii.contiguous(), jjc = jj.contiguous();
  torch::Tensor g = corr_grad.to(torch::kFloat).contiguous();
  const int dt = wv_dtype_of(f1, "corr");
  const int B = co.size(0), E = co.size(1), PH = co.size(3), PW = co.size(4);
  const int n1 = f1.size(1), C = f1.size(2), n2 = f2.size(1), H2 = f2.size(3), W2 = f2.size(4);
  const int T = 2 * radius + 1;
  TORCH_CHECK(g.numel() == (int64_t)B * E * T * T * PH * PW, "corr backward: gradient has the wrong number of elements");
  torch::Tensor g1 = torch::zeros(f1.sizes(), f1.options().dtype(torch::kFloat));
  torch::Tensor g2 = torch::zeros(f2.sizes(), f2.options().dtype(torch::kFloat));
  check_rc(wv_corr_backward(f1.data_ptr(), f2.data_ptr(), co.data_ptr<float>(), iic.data_ptr<int64_t>(), jjc.data_ptr<int64_t>(),
                            g.data_ptr<float>(), g1.data_ptr<float>(), g2.data_ptr<float>(), dt, B, E, C, PH, PW, n1, n2, H2,
                            W2, radius, status_word(f1.device()).data_ptr<int>(), current_stream()),
           "cuda_corr.backward");
  return {g1.to(in_dtype), g2.to(in_dtype)};
}

std::vector<torch::Tensor> patchify_forward(torch::Tensor net, torch::Tensor coords, int radius) {
  check_cuda(net, "net"); check_cuda(coords, "coords");
  TORCH_CHECK(net.dim() == 4 && coords.dim() == 3 && coords.size(2) == 2, "patchify: net must be [B,C,H,W], coords [B,M,2]");
  TORCH_CHECK(coords.scalar_type() == torch::kFloat, "patchify: coords must be float32");
  c10::cuda::CUDAGuard guard(net.device());
  const auto in_dtype = net.scalar_type();
  torch::Tensor nt = as_map(net), co = coords.contiguous();
  const int dt = wv_dtype_of(nt, "patchify");
  const int B = co.size(0), M = co.size(1), C = nt.size(1), H = nt.size(2), W = nt.size(3), D = 2 * radius + 2;
  torch::Tensor patches = torch::empty({B, M, C, D, D}, nt.options());
  check_rc(wv_patchify_forward(nt.data_ptr(), co.data_ptr<float>(), patches.data_ptr(), dt, B, M, C, H, W, radius, current_stream()),
           "cuda_corr.patchify_forward");
  if (in_dtype == torch::kDouble) patches = patches.to(torch::kDouble);
  return {patches};
}

std::vector<torch::Tensor> patchify_backward(torch::Tensor net, torch::Tensor coords, torch::Tensor gradient, int radius) {
  check_cuda(net, "net"); check_cuda(coords, "coords"); check_cuda(gradient, "gradient");
  c10::cuda::CUDAGuard guard(net.device());
  const auto in_dtype = net.scalar_type();
  torch::Tensor co = coords.contiguous();
  torch::Tensor g = (gradient.scalar_type() == torch::kDouble ? gradient.to(torch::kFloat) : gradient).contiguous();
  const int dt = wv_dtype_of(g, "patchify");
  const int B = co.size(0), M = co.size(1), C = net.size(1), H = net.size(2), W = net.size(3);
  torch::Tensor ng = torch::zeros(net.sizes(), net.options().dtype(torch::kFloat));
  check_rc(wv_patchify_backward(g.data_ptr(), co.data_ptr<float>(), ng.data_ptr<float>(), dt, B, M, C, H, W, radius, current_stream()),
           "cuda_corr.patchify_backward");
  return {ng.to(in_dtype)};
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("forward", &corr_forward, "CORR forward");
  m.def("backward", &corr_backward, "CORR backward");
  m.def("forward_stacked", &corr_forward_stacked, "both pyramid levels, stacked (DPVO.corr)");  // beyond the reference's table
  m.def("patchify_forward", &patchify_forward, "PATCHIFY forward");
  m.def("patchify_backward", &patchify_backward, "PATCHIFY backward");
}
