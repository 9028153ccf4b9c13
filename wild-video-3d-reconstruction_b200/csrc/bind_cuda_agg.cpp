// torch extension `cuda_agg`: the update operator's SoftAgg (dpvo/blocks.py:31-48).  The reference has no extension of its own
// here -- it calls torch.unique and torch_scatter (blocks.py:41-43); this module is what blocks.SoftAgg of this package binds.
#include "bind_common.h"

using namespace wvbind;

namespace {

// `_, jx = torch.unique(ix, return_inverse=True)` plus the member lists the aggregation kernel walks.
// returns (jx [E] int64, order [E] int32, offsets [E+1] int32, n_groups [2] int32 on the device (G, status of this call), G as a
// host int).  key_min / key_max bound the keys (inclusive); pass key_max < key_min to have them measured here (one extra
// reduction + sync).  sync = false skips the host read of G (returned as -1; bad keys then go unreported): for callers that
// size y by a bound on G instead, so that nothing in the update operator waits for the device.
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, int64_t> group_plan(torch::Tensor ix, int64_t key_min,
                                                                                          int64_t key_max, bool sync) {
  check_cuda(ix, "ix");
  TORCH_CHECK(ix.scalar_type() == torch::kLong && ix.dim() == 1, "group_plan: ix must be a 1-d int64 tensor");
  c10::cuda::CUDAGuard guard(ix.device());
  torch::Tensor keys = ix.contiguous();
  const int64_t E = keys.numel();
  auto i32 = torch::dtype(torch::kInt32).device(ix.device());
  torch::Tensor jx = torch::empty({E}, keys.options()), order = torch::empty({E}, i32), offsets = torch::zeros({E + 1}, i32),
                ng = torch::zeros({2}, i32);
  if (E == 0) return {jx, order, offsets, ng, 0};
  if (key_max < key_min) {
    auto mm = torch::aminmax(keys);
    key_min = std::get<0>(mm).item<int64_t>();
    key_max = std::get<1>(mm).item<int64_t>();
  }
  const int64_t range = key_max - key_min + 1;
  TORCH_CHECK(range > 0 && range <= (1LL << 28), "group_plan: key range ", range, " outside (0, 2^28]");
  const size_t wsb = wv_group_plan_workspace_bytes(E, range);
  TORCH_CHECK(wsb > 0, "group_plan: unsupported size");
  torch::Tensor ws = torch::empty({(int64_t)wsb}, torch::dtype(torch::kUInt8).device(ix.device()));
  check_rc(wv_group_plan(keys.data_ptr<int64_t>(), E, key_min, range, jx.data_ptr<int64_t>(), order.data_ptr<int>(),
                         offsets.data_ptr<int>(), ng.data_ptr<int>(), ws.data_ptr(), wsb, ng.data_ptr<int>() + 1, current_stream()),
           "cuda_agg.group_plan");
  if (!sync) return {jx, order, offsets, ng, -1};
  // G sizes the output, as the length of torch.unique's first result does in the reference: one host read (G and the status)
  int gs[2] = {0, 0};
  check_rc(wv_words_read(ng.data_ptr<int>(), 2, gs, current_stream()), "cuda_agg.group_plan");
  TORCH_CHECK_INDEX(!(gs[1] & WV_STATUS_BAD_INDEX), "cuda_agg.group_plan: a key lies outside [key_min, key_max]");
  return {jx, order, offsets, ng, (int64_t)gs[0]};
}

// y [B, G, D] = scatter_sum(fx * scatter_softmax(gx, jx, dim=1), jx, dim=1); G may be a bound on the group count (extra rows zero)
torch::Tensor soft_agg(torch::Tensor fx, torch::Tensor gx, torch::Tensor order, torch::Tensor offsets, torch::Tensor n_groups, int64_t G) {
  check_cuda(fx, "fx"); check_cuda(gx, "gx"); check_cuda(order, "order"); check_cuda(offsets, "offsets"); check_cuda(n_groups, "n_groups");
  TORCH_CHECK(fx.dim() == 3 && gx.sizes() == fx.sizes() && fx.scalar_type() == gx.scalar_type(), "soft_agg: fx and gx must both be [B, E, D]");
  TORCH_CHECK(order.scalar_type() == torch::kInt32 && offsets.scalar_type() == torch::kInt32 && n_groups.scalar_type() == torch::kInt32,
              "soft_agg: plan tensors must be int32");
  c10::cuda::CUDAGuard guard(fx.device());
  torch::Tensor f = fx.contiguous(), g = gx.contiguous();
  const int64_t B = f.size(0), E = f.size(1), D = f.size(2);
  TORCH_CHECK(order.numel() == E && offsets.numel() == E + 1 && G >= 0, "soft_agg: plan does not match the edge count");
  const int dt = wv_dtype_of(f, "soft_agg");
  torch::Tensor y = torch::empty({B, G, D}, f.options());
  check_rc(wv_soft_agg_forward(f.data_ptr(), g.data_ptr(), order.contiguous().data_ptr<int>(), offsets.contiguous().data_ptr<int>(),
                               n_groups.data_ptr<int>(), y.data_ptr(), dt, (int)B, E, (int)D, G, current_stream()),
           "cuda_agg.soft_agg");
  return y;
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("group_plan", &group_plan, "unique(ix, return_inverse=True) + member lists", py::arg("ix"), py::arg("key_min") = 0,
        py::arg("key_max") = -1, py::arg("sync") = true);
  m.def("soft_agg", &soft_agg, "fused scatter_softmax + scatter_sum");
}
