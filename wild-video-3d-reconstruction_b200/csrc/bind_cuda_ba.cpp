// torch extension `cuda_ba`: same module name and m.def table as the reference's dpvo/fastba/ba.cpp:236-241
// (forward / neighbors / reproject / solve_system), so dpvo/fastba/ba.py:2 and
// dpvo/loop_closure/optim_utils.py:1 import it unchanged.  `forward` additionally takes two OPTIONAL trailing
// arguments (patches_est, mu) for the depth-prior term of dpvo/ba.py:151-159; the 12-argument call is the
// reference behaviour.
#include "bind_common.h"

using namespace wvbind;

namespace {

torch::Tensor workspace(const torch::Device& dev, size_t bytes) {
  // grown on demand, kept per device: the plan never needs it cleared (ba_plan.cuh)
  static std::map<int, torch::Tensor> ws;
  auto it = ws.find(dev.index());
  if (it == ws.end() || (size_t)it->second.numel() < bytes) {
    ws[dev.index()] = torch::empty({(int64_t)(bytes + (bytes >> 2) + 4096)}, torch::dtype(torch::kUInt8).device(dev));
    it = ws.find(dev.index());
  }
  return it->second;
}

// The tracker calls neighbors(kk, jj) and then BA(..., ii, jj, kk, ...) on the same edge tensors inside one update()
// (dpvo/net.py:81, dpvo/dpvo.py:733).  neighbors leaves its edge plan in the per-device workspace; if BA then sees the same
// storage (pointer, length, version counter), workspace and stream, it reuses the plan instead of rebuilding it.
constexpr int64_t kPlanGroups = (int64_t)1 << 22;  // ids (patch indices) must lie below 4M for the shared plan
struct PlanNote {
  const void* kk = nullptr; const void* jj = nullptr; int64_t E = -1; uint32_t vkk = 0, vjj = 0;
  const void* ws = nullptr; void* stream = nullptr;
};
PlanNote& plan_note(int dev) {
  static std::map<int, PlanNote> notes;
  return notes[dev];
}

void check_f32c(const torch::Tensor& t, const char* name) {
  check_cuda(t, name);
  TORCH_CHECK(t.scalar_type() == torch::kFloat, "cuda_ba: ", name, " must be float32");
  TORCH_CHECK(t.is_contiguous(), "cuda_ba: ", name, " must be contiguous");
}

torch::Tensor idx(const torch::Tensor& t, const char* name) {
  check_cuda(t, name);
  TORCH_CHECK(t.scalar_type() == torch::kLong, "cuda_ba: ", name, " must be int64");
  return t.contiguous().view({-1});
}

// ---- BA with more than 16 free poses (global BA): build / solve / dz / retract pieces over the C ABI's wv_ba_global_* -------
struct BAIn {
  torch::Tensor tg, wt, lm, ii, jj, kk, est;
  const float* estp = nullptr;
  int P = 0; int64_t E = 0, n_poses = 0, n_patches = 0;
};

BAIn ba_inputs(const torch::Tensor& poses, const torch::Tensor& patches, const torch::Tensor& intrinsics, const torch::Tensor& target,
               const torch::Tensor& weight, const torch::Tensor& lmbda, const torch::Tensor& ii, const torch::Tensor& jj,
               const torch::Tensor& kk, int t0, int t1, const c10::optional<torch::Tensor>& patches_est) {
  check_f32c(poses, "poses"); check_f32c(patches, "patches"); check_f32c(intrinsics, "intrinsics");
  TORCH_CHECK(patches.dim() >= 4, "cuda_ba: patches must be [..., 3, P, P]");
  BAIn in;
  in.P = patches.size(-1);
  in.n_poses = poses.numel() / 7; in.n_patches = patches.numel() / (3 * in.P * in.P);
  in.tg = target.to(torch::kFloat).contiguous().view({-1, 2}); in.wt = weight.to(torch::kFloat).contiguous().view({-1, 2});
  in.lm = lmbda.to(torch::kFloat).contiguous().view({-1});
  check_cuda(in.tg, "target"); check_cuda(in.wt, "weight"); check_cuda(in.lm, "lmbda");
  in.ii = idx(ii, "ii"); in.jj = idx(jj, "jj"); in.kk = idx(kk, "kk");
  in.E = in.ii.numel();
  TORCH_CHECK(in.jj.numel() == in.E && in.kk.numel() == in.E && in.tg.size(0) == in.E && in.wt.size(0) == in.E,
              "cuda_ba: inconsistent edge count");
  TORCH_CHECK(t0 >= 0 && t1 >= t0 && t1 <= in.n_poses, "cuda_ba: bad pose window [t0, t1)");
  if (patches_est.has_value() && patches_est->defined()) {
    in.est = patches_est->to(torch::kFloat).contiguous();
    TORCH_CHECK(in.est.numel() == patches.numel() && in.est.is_cuda(), "cuda_ba: patches_est must match patches");
    in.estp = in.est.data_ptr<float>();
  }
  return in;
}

torch::Tensor global_ws(const torch::Device& dev, const BAIn& in) {
  plan_note(dev.index()) = PlanNote();  // the shared workspace now holds this call's plan
  return workspace(dev, wv_ba_global_workspace_bytes(in.E, in.n_patches));
}

// partial reduced system of this shard: returns one flat buffer [6N*6N + 6N] = [S (row-major, lower triangle) | y], so that a
// multi-GPU caller sums it with ONE allreduce
torch::Tensor global_build(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor target,
                           torch::Tensor weight, torch::Tensor lmbda, torch::Tensor ii, torch::Tensor jj, torch::Tensor kk, int t0,
                           int t1, int shard, int nshards, bool build_plan, c10::optional<torch::Tensor> patches_est, double mu) {
  c10::cuda::CUDAGuard guard(poses.device());
  BAIn in = ba_inputs(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, patches_est);
  const int64_t n6 = 6 * (int64_t)(t1 - t0);
  torch::Tensor Sy = torch::empty({n6 * n6 + n6}, poses.options());
  if (in.E == 0) return Sy.zero_();
  torch::Tensor ws = build_plan ? global_ws(poses.device(), in) : workspace(poses.device(), wv_ba_global_workspace_bytes(in.E, in.n_patches));
  check_rc(wv_ba_global_build(poses.data_ptr<float>(), patches.data_ptr<float>(), intrinsics.data_ptr<float>(), in.tg.data_ptr<float>(),
                              in.wt.data_ptr<float>(), in.lm.data_ptr<float>(), in.ii.data_ptr<int64_t>(), in.jj.data_ptr<int64_t>(),
                              in.kk.data_ptr<int64_t>(), in.E, in.n_poses, in.n_patches, in.P, t0, t1, in.estp, (float)mu, shard,
                              nshards, build_plan ? 1 : 0, Sy.data_ptr<float>(), Sy.data_ptr<float>() + n6 * n6, ws.data_ptr(),
                              (size_t)ws.numel(), status_word(poses.device()).data_ptr<int>(), current_stream()),
           "cuda_ba.global_build");
  return Sy;
}

// widest frame span of a patch's observations among the free poses (one host sync): decides band vs dense.  Builds the edge plan.
int64_t global_span(torch::Tensor poses, torch::Tensor patches, torch::Tensor ii, torch::Tensor jj, torch::Tensor kk, int t0, int t1,
                    bool build_plan) {
  check_f32c(poses, "poses"); check_f32c(patches, "patches");
  c10::cuda::CUDAGuard guard(poses.device());
  torch::Tensor iic = idx(ii, "ii"), jjc = idx(jj, "jj"), kkc = idx(kk, "kk");
  const int64_t E = iic.numel();
  if (E == 0) return 0;
  const int P = patches.size(-1);
  const int64_t n_poses = poses.numel() / 7, n_patches = patches.numel() / (3 * P * P);
  if (build_plan) plan_note(poses.device().index()) = PlanNote();  // the shared workspace now holds this call's plan
  torch::Tensor ws = workspace(poses.device(), wv_ba_global_workspace_bytes(E, n_patches));
  torch::Tensor out = torch::zeros({1}, torch::dtype(torch::kInt32).device(poses.device()));
  check_rc(wv_ba_global_span(iic.data_ptr<int64_t>(), jjc.data_ptr<int64_t>(), kkc.data_ptr<int64_t>(), E, n_poses, n_patches, t0, t1,
                             build_plan ? 1 : 0, out.data_ptr<int>(), ws.data_ptr(), (size_t)ws.numel(),
                             status_word(poses.device()).data_ptr<int>(), current_stream()),
           "cuda_ba.global_span");
  return out.item<int>();
}

// half-width (frames) of the band path for a given span, or 0 when the dense path must be taken
int64_t band_block(int64_t span) {
  const int64_t bs = std::max<int64_t>(span, 2);
  return bs <= wv_ba_band_max_bs() && !env_flag("WVB200_GLOBAL_DENSE") ? bs : 0;
}

// this shard's partial band buffer [band | y] (csrc/ba_band.cu); the edge plan must be in the workspace (global_span built it)
torch::Tensor global_build_band(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor target,
                                torch::Tensor weight, torch::Tensor lmbda, torch::Tensor ii, torch::Tensor jj, torch::Tensor kk, int t0,
                                int t1, int shard, int nshards, int64_t bs, c10::optional<torch::Tensor> patches_est, double mu) {
  c10::cuda::CUDAGuard guard(poses.device());
  BAIn in = ba_inputs(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, patches_est);
  const int64_t n6 = 6 * (int64_t)(t1 - t0);
  const int64_t nb = ((int64_t)wv_ba_band_floats((int)n6, (int)bs) + 3) / 4 * 4;   // y starts 16-byte aligned
  torch::Tensor Sy = torch::empty({nb + n6}, poses.options());
  if (in.E == 0) return Sy.zero_();
  torch::Tensor ws = workspace(poses.device(), wv_ba_global_workspace_bytes(in.E, in.n_patches));
  check_rc(wv_ba_global_build_band(poses.data_ptr<float>(), patches.data_ptr<float>(), intrinsics.data_ptr<float>(), in.tg.data_ptr<float>(),
                                   in.wt.data_ptr<float>(), in.lm.data_ptr<float>(), in.ii.data_ptr<int64_t>(), in.jj.data_ptr<int64_t>(),
                                   in.kk.data_ptr<int64_t>(), in.E, in.n_poses, in.n_patches, in.P, t0, t1, in.estp, (float)mu, shard,
                                   nshards, 0, (int)bs, Sy.data_ptr<float>(), Sy.data_ptr<float>() + nb, ws.data_ptr(), (size_t)ws.numel(),
                                   status_word(poses.device()).data_ptr<int>(), current_stream()),
           "cuda_ba.global_build_band");
  return Sy;
}

// damping + band Cholesky solve of the band buffer: dX [N, 6]
torch::Tensor global_solve_band(torch::Tensor Sy, int n_free, int64_t bs) {
  c10::cuda::CUDAGuard guard(Sy.device());
  const int64_t n6 = 6 * (int64_t)n_free;
  const int64_t nb = ((int64_t)wv_ba_band_floats((int)n6, (int)bs) + 3) / 4 * 4;
  TORCH_CHECK(Sy.is_cuda() && Sy.scalar_type() == torch::kFloat && Sy.is_contiguous() && Sy.numel() == nb + n6,
              "cuda_ba.global_solve_band: expected the flat [band | y] buffer of global_build_band");
  torch::Tensor dX = torch::empty({n_free, 6}, Sy.options());
  if (n6 == 0) return dX;
  torch::Tensor F = torch::empty({(int64_t)wv_ba_band_factor_floats((int)n6, (int)bs)}, Sy.options());
  check_rc(wv_ba_band_solve(Sy.data_ptr<float>(), Sy.data_ptr<float>() + nb, (int)n6, (int)bs, 1, F.data_ptr<float>(), dX.data_ptr<float>(),
                            status_word(Sy.device()).data_ptr<int>(), current_stream()),
           "cuda_ba.global_solve_band");
  return dX;
}

// damping + the reference's own dense solve (ba_cuda.cu:517-523): returns dX [N, 6]
torch::Tensor global_solve(torch::Tensor Sy, int n_free) {
  c10::cuda::CUDAGuard guard(Sy.device());
  const int64_t n6 = 6 * (int64_t)n_free;
  TORCH_CHECK(Sy.is_cuda() && Sy.scalar_type() == torch::kFloat && Sy.is_contiguous() && Sy.numel() == n6 * n6 + n6,
              "cuda_ba.global_solve: expected the flat [S | y] buffer of global_build");
  if (n6 == 0) return torch::zeros({0, 6}, Sy.options());
  check_rc(wv_ba_global_damp(Sy.data_ptr<float>(), (int)n6, current_stream()), "cuda_ba.global_solve");
  torch::Tensor S = Sy.narrow(0, 0, n6 * n6).view({n6, n6}), y = Sy.narrow(0, n6 * n6, n6).view({n6, 1});
  auto [L, info] = at::linalg_cholesky_ex(S, /*upper=*/false, /*check_errors=*/false);
  TORCH_CHECK(info.item<int>() == 0, "cuda_ba.forward: linalg.cholesky: the reduced camera system is not positive-definite");
  return at::cholesky_solve(y, L, /*upper=*/false).view({n_free, 6}).contiguous();
}

// this shard's depth updates (zeros elsewhere; a multi-GPU caller sums over ranks): dZ [min(E, n_patches)]
torch::Tensor global_dz(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor target,
                        torch::Tensor weight, torch::Tensor lmbda, torch::Tensor ii, torch::Tensor jj, torch::Tensor kk, int t0, int t1,
                        int shard, int nshards, torch::Tensor dX, c10::optional<torch::Tensor> patches_est, double mu) {
  c10::cuda::CUDAGuard guard(poses.device());
  BAIn in = ba_inputs(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, patches_est);
  TORCH_CHECK(dX.is_cuda() && dX.scalar_type() == torch::kFloat && dX.is_contiguous() && dX.numel() == 6 * (int64_t)(t1 - t0),
              "cuda_ba.global_dz: dX must be float32 [t1 - t0, 6]");
  torch::Tensor dZ = torch::empty({std::min(in.E, in.n_patches)}, poses.options());
  if (in.E == 0) return dZ;
  torch::Tensor ws = workspace(poses.device(), wv_ba_global_workspace_bytes(in.E, in.n_patches));
  check_rc(wv_ba_global_dz(poses.data_ptr<float>(), patches.data_ptr<float>(), intrinsics.data_ptr<float>(), in.tg.data_ptr<float>(),
                           in.wt.data_ptr<float>(), in.lm.data_ptr<float>(), in.ii.data_ptr<int64_t>(), in.jj.data_ptr<int64_t>(), in.E,
                           in.n_poses, in.n_patches, in.P, t0, t1, in.estp, (float)mu, shard, nshards, dX.data_ptr<float>(),
                           dZ.data_ptr<float>(), ws.data_ptr(), (size_t)ws.numel(), status_word(poses.device()).data_ptr<int>(),
                           current_stream()),
           "cuda_ba.global_dz");
  return dZ;
}

void global_retract(torch::Tensor poses, torch::Tensor patches, int64_t E, int t0, int t1, torch::Tensor dX, torch::Tensor dZ) {
  check_f32c(poses, "poses"); check_f32c(patches, "patches");
  c10::cuda::CUDAGuard guard(poses.device());
  const int P = patches.size(-1);
  const int64_t n_poses = poses.numel() / 7, n_patches = patches.numel() / (3 * P * P);
  TORCH_CHECK(dX.is_cuda() && dX.scalar_type() == torch::kFloat && dX.is_contiguous() && dX.numel() == 6 * (int64_t)(t1 - t0) &&
                  dZ.is_cuda() && dZ.scalar_type() == torch::kFloat && dZ.is_contiguous() && dZ.numel() == std::min(E, n_patches),
              "cuda_ba.global_retract: dX / dZ do not match this problem");
  if (E == 0) return;
  torch::Tensor ws = workspace(poses.device(), wv_ba_global_workspace_bytes(E, n_patches));
  check_rc(wv_ba_global_retract(poses.data_ptr<float>(), patches.data_ptr<float>(), E, n_poses, n_patches, P, t0, t1,
                                dX.data_ptr<float>(), dZ.data_ptr<float>(), ws.data_ptr(), (size_t)ws.numel(), current_stream()),
           "cuda_ba.global_retract");
}

std::vector<torch::Tensor> ba_forward(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor target,
                                      torch::Tensor weight, torch::Tensor lmbda, torch::Tensor ii, torch::Tensor jj,
                                      torch::Tensor kk, int t0, int t1, int iterations,
                                      c10::optional<torch::Tensor> patches_est, double mu) {
  check_f32c(poses, "poses"); check_f32c(patches, "patches"); check_f32c(intrinsics, "intrinsics");
  TORCH_CHECK(patches.dim() >= 4, "cuda_ba: patches must be [..., 3, P, P]");
  c10::cuda::CUDAGuard guard(poses.device());
  const int P = patches.size(-1);
  const int64_t n_poses = poses.numel() / 7, n_patches = patches.numel() / (3 * P * P);
  torch::Tensor tg = target.to(torch::kFloat).contiguous().view({-1, 2}), wt = weight.to(torch::kFloat).contiguous().view({-1, 2});
  torch::Tensor lm = lmbda.to(torch::kFloat).contiguous().view({-1});
  check_cuda(tg, "target"); check_cuda(wt, "weight"); check_cuda(lm, "lmbda");
  torch::Tensor iic = idx(ii, "ii"), jjc = idx(jj, "jj"), kkc = idx(kk, "kk");
  const int64_t E = iic.numel();
  TORCH_CHECK(jjc.numel() == E && kkc.numel() == E && tg.size(0) == E && wt.size(0) == E, "cuda_ba: inconsistent edge count");
  TORCH_CHECK(t0 >= 0 && t1 >= t0 && t1 <= n_poses, "cuda_ba: bad pose window [t0, t1)");
  const float* est = nullptr;
  torch::Tensor estc;
  if (patches_est.has_value() && patches_est->defined()) {
    estc = patches_est->to(torch::kFloat).contiguous();
    TORCH_CHECK(estc.numel() == patches.numel() && estc.is_cuda(), "cuda_ba: patches_est must match patches");
    est = estc.data_ptr<float>();
  }
  if (E == 0 || iterations <= 0) return {};
  if (t1 - t0 > 16) {  // global BA: reduced system in HBM -- banded (band Cholesky, csrc/ba_band.cu) when the edge list allows, else dense
    const int64_t bs = band_block(global_span(poses, patches, iic, jjc, kkc, t0, t1, true));
    for (int it = 0; it < iterations; it++) {
      torch::Tensor Sy = bs ? global_build_band(poses, patches, intrinsics, tg, wt, lm, iic, jjc, kkc, t0, t1, 0, 1, bs, patches_est, mu)
                            : global_build(poses, patches, intrinsics, tg, wt, lm, iic, jjc, kkc, t0, t1, 0, 1, false, patches_est, mu);
      raise_on_status(poses.device(), "cuda_ba.forward");  // bad index / over-dense patch: say so before the solve trips over it
      torch::Tensor dX = bs ? global_solve_band(Sy, t1 - t0, bs) : global_solve(Sy, t1 - t0);
      Sy = torch::Tensor();
      torch::Tensor dZ = global_dz(poses, patches, intrinsics, tg, wt, lm, iic, jjc, kkc, t0, t1, 0, 1, dX, patches_est, mu);
      global_retract(poses, patches, E, t0, t1, dX, dZ);
    }
    raise_on_status(poses.device(), "cuda_ba.forward");
    return {};
  }
  PlanNote& note = plan_note(poses.device().index());
  const size_t wsb_shared = wv_ba_workspace_bytes(E, kPlanGroups, t1 - t0);
  torch::Tensor ws = workspace(poses.device(), std::max(wv_ba_workspace_bytes(E, n_patches, t1 - t0), note.E == E ? wsb_shared : (size_t)0));
  const bool planned = !env_flag("WVB200_NO_PLAN_CACHE") && note.E == E && note.kk == kkc.data_ptr() && note.jj == jjc.data_ptr() &&
                       note.vkk == kkc._version() && note.vjj == jjc._version() && note.ws == ws.data_ptr() &&
                       note.stream == (void*)current_stream() && n_patches <= kPlanGroups && (size_t)ws.numel() >= wsb_shared;
  auto own_plan = [&]() {
    check_rc(wv_ba_forward(poses.data_ptr<float>(), patches.data_ptr<float>(), intrinsics.data_ptr<float>(), tg.data_ptr<float>(),
                           wt.data_ptr<float>(), lm.data_ptr<float>(), iic.data_ptr<int64_t>(), jjc.data_ptr<int64_t>(),
                           kkc.data_ptr<int64_t>(), E, n_poses, n_patches, P, t0, t1, iterations, est, (float)mu, ws.data_ptr(),
                           (size_t)ws.numel(), status_word(poses.device()).data_ptr<int>(), current_stream()),
             "cuda_ba.forward");
  };
  if (planned) {
    check_rc(wv_ba_forward_planned(poses.data_ptr<float>(), patches.data_ptr<float>(), intrinsics.data_ptr<float>(),
                                   tg.data_ptr<float>(), wt.data_ptr<float>(), lm.data_ptr<float>(), iic.data_ptr<int64_t>(),
                                   jjc.data_ptr<int64_t>(), kkc.data_ptr<int64_t>(), E, n_poses, n_patches, P, t0, t1, iterations, est,
                                   (float)mu, ws.data_ptr(), (size_t)ws.numel(), kPlanGroups, status_word(poses.device()).data_ptr<int>(),
                                   current_stream()),
             "cuda_ba.forward");
    note = PlanNote();  // one-shot: the next BA without a neighbors() before it builds its own plan
  } else {
    note = PlanNote();  // this call rebuilds the plan with its own layout
    own_plan();
  }
  // The reference raises from torch::linalg::cholesky inside this call (dpvo/dpvo.py:733 lets it propagate);
  // reading the status word costs one host sync.  WVB200_ASYNC=1 skips it (capture-safe, errors stay sticky).
  if (!env_flag("WVB200_ASYNC")) {
    int s = read_status(poses.device(), "cuda_ba.forward");
    if (planned && (s & WV_STATUS_STALE_PLAN)) {
      // the plan neighbors() left behind was not for this kk (a caching-allocator address reused by another tensor): the kernel
      // stopped before applying anything, so the call is simply repeated with its own plan
      own_plan();
      s = (s & ~WV_STATUS_STALE_PLAN) | read_status(poses.device(), "cuda_ba.forward");
    }
    raise_for_status(s, "cuda_ba.forward");
  }
  return {};
}

std::vector<torch::Tensor> neighbors(torch::Tensor ii, torch::Tensor jj) {
  torch::Tensor iic = idx(ii, "ii"), jjc = idx(jj, "jj");
  TORCH_CHECK(iic.numel() == jjc.numel(), "neighbors: ii and jj must have the same length");
  c10::cuda::CUDAGuard guard(iic.device());
  const int64_t E = iic.numel();
  torch::Tensor ix = torch::empty({E}, iic.options()), jx = torch::empty({E}, iic.options());
  if (E == 0) return {ix, jx};
  const int64_t n_groups = kPlanGroups;
  // room for a following windowed BA on the same plan (<= 16 free poses) -- only at the tracker's edge counts: the global-BA list
  // (E ~ 1e7, also sent through network.update -> neighbors) would pin gigabytes for a plan the global path cannot use anyway
  const size_t wsb = wv_ba_workspace_bytes(E, n_groups, E <= ((int64_t)1 << 20) ? 16 : 0);
  torch::Tensor ws = workspace(iic.device(), wsb);
  PlanNote& note = plan_note(iic.device().index());
  note = PlanNote();
  check_rc(wv_neighbors(iic.data_ptr<int64_t>(), jjc.data_ptr<int64_t>(), E, n_groups, ix.data_ptr<int64_t>(),
                        jx.data_ptr<int64_t>(), ws.data_ptr(), (size_t)ws.numel(), status_word(iic.device()).data_ptr<int>(),
                        current_stream()),
           "cuda_ba.neighbors");
  note.kk = iic.data_ptr(); note.jj = jjc.data_ptr(); note.E = E; note.vkk = iic._version(); note.vjj = jjc._version();
  note.ws = ws.data_ptr(); note.stream = (void*)current_stream();
  if (env_flag("WVB200_CHECK_INDICES")) raise_on_status(iic.device(), "cuda_ba.neighbors");
  return {ix, jx};
}

torch::Tensor reproject(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor ii, torch::Tensor jj,
                        torch::Tensor kk) {
  check_cuda(poses, "poses"); check_cuda(patches, "patches"); check_cuda(intrinsics, "intrinsics");
  c10::cuda::CUDAGuard guard(poses.device());
  torch::Tensor po = poses.to(torch::kFloat).contiguous(), pa = patches.to(torch::kFloat).contiguous(),
                in = intrinsics.to(torch::kFloat).contiguous();
  const int P = pa.size(-1);
  const int64_t n_poses = po.numel() / 7, n_patches = pa.numel() / (3 * P * P);
  torch::Tensor iic = idx(ii, "ii"), jjc = idx(jj, "jj"), kkc = idx(kk, "kk");
  const int64_t E = iic.numel();
  torch::Tensor coords = torch::empty({1, E, 2, P, P}, po.options());
  check_rc(wv_ba_reproject(po.data_ptr<float>(), pa.data_ptr<float>(), in.data_ptr<float>(), iic.data_ptr<int64_t>(),
                           jjc.data_ptr<int64_t>(), kkc.data_ptr<int64_t>(), E, n_poses, n_patches, P, coords.data_ptr<float>(),
                           status_word(po.device()).data_ptr<int>(), current_stream()),
           "cuda_ba.reproject");
  return coords;
}

// ---- fused projective ops (dpvo/projective_ops.py over lietorch), SURVEY 8f N1 / N4 ---------------------------------------
struct PopsIn { torch::Tensor po, pa, in, ii, jj, kk; int P; int64_t E, n_poses, n_patches; };

PopsIn pops_inputs(const torch::Tensor& poses, const torch::Tensor& patches, const torch::Tensor& intrinsics, const torch::Tensor& ii,
                   const torch::Tensor& jj, const torch::Tensor& kk) {
  check_cuda(poses, "poses"); check_cuda(patches, "patches"); check_cuda(intrinsics, "intrinsics");
  PopsIn x;
  x.po = poses.to(torch::kFloat).contiguous(); x.pa = patches.to(torch::kFloat).contiguous(); x.in = intrinsics.to(torch::kFloat).contiguous();
  TORCH_CHECK(x.pa.dim() >= 4, "projective ops: patches must be [..., 3, P, P]");
  x.P = x.pa.size(-1);
  x.n_poses = x.po.numel() / 7; x.n_patches = x.pa.numel() / (3 * x.P * x.P);
  TORCH_CHECK(x.in.numel() == 4 * x.n_poses, "projective ops: intrinsics must have one row per pose");
  x.ii = idx(ii, "ii"); x.jj = idx(jj, "jj"); x.kk = idx(kk, "kk");
  x.E = x.ii.numel();
  TORCH_CHECK(x.jj.numel() == x.E && x.kk.numel() == x.E, "projective ops: ii, jj, kk must have the same length");
  return x;
}

std::vector<torch::Tensor> pops_transform(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor ii,
                                          torch::Tensor jj, torch::Tensor kk, bool depth, bool valid, bool tonly) {
  c10::cuda::CUDAGuard guard(poses.device());
  PopsIn x = pops_inputs(poses, patches, intrinsics, ii, jj, kk);
  torch::Tensor coords = torch::empty({1, x.E, x.P, x.P, depth ? 3 : 2}, x.po.options());
  torch::Tensor val = valid ? torch::empty({1, x.E, x.P, x.P}, x.po.options()) : torch::Tensor();
  check_rc(wv_pops_transform(x.po.data_ptr<float>(), x.pa.data_ptr<float>(), x.in.data_ptr<float>(), x.ii.data_ptr<int64_t>(),
                             x.jj.data_ptr<int64_t>(), x.kk.data_ptr<int64_t>(), x.E, x.n_poses, x.n_patches, x.P, tonly, depth,
                             coords.data_ptr<float>(), valid ? val.data_ptr<float>() : nullptr,
                             status_word(x.po.device()).data_ptr<int>(), current_stream()),
           "cuda_ba.transform");
  if (env_flag("WVB200_CHECK_INDICES")) raise_on_status(x.po.device(), "cuda_ba.transform");
  if (valid) return {coords, val};
  return {coords};
}

// transform(..., jacobian=True): [x1, validc, Ji, Jj, Jz] (projective_ops.py:71-99)
std::vector<torch::Tensor> pops_transform_jacobian(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor ii,
                                                   torch::Tensor jj, torch::Tensor kk, bool depth, bool tonly) {
  c10::cuda::CUDAGuard guard(poses.device());
  PopsIn x = pops_inputs(poses, patches, intrinsics, ii, jj, kk);
  torch::Tensor coords = torch::empty({1, x.E, x.P, x.P, depth ? 3 : 2}, x.po.options());
  torch::Tensor Ji = torch::empty({1, x.E, 2, 6}, x.po.options()), Jj = torch::empty({1, x.E, 2, 6}, x.po.options());
  torch::Tensor Jz = torch::empty({1, x.E, 2, 1}, x.po.options()), val = torch::empty({1, x.E}, x.po.options());
  int* st = status_word(x.po.device()).data_ptr<int>();
  check_rc(wv_pops_transform(x.po.data_ptr<float>(), x.pa.data_ptr<float>(), x.in.data_ptr<float>(), x.ii.data_ptr<int64_t>(),
                             x.jj.data_ptr<int64_t>(), x.kk.data_ptr<int64_t>(), x.E, x.n_poses, x.n_patches, x.P, tonly, depth,
                             coords.data_ptr<float>(), nullptr, st, current_stream()),
           "cuda_ba.transform_jacobian");
  check_rc(wv_pops_jacobian(x.po.data_ptr<float>(), x.pa.data_ptr<float>(), x.in.data_ptr<float>(), x.ii.data_ptr<int64_t>(),
                            x.jj.data_ptr<int64_t>(), x.kk.data_ptr<int64_t>(), x.E, x.n_poses, x.n_patches, x.P, tonly,
                            Ji.data_ptr<float>(), Jj.data_ptr<float>(), Jz.data_ptr<float>(), val.data_ptr<float>(), st, current_stream()),
           "cuda_ba.transform_jacobian");
  if (env_flag("WVB200_CHECK_INDICES")) raise_on_status(x.po.device(), "cuda_ba.transform_jacobian");
  return {coords, val, Ji, Jj, Jz};
}

torch::Tensor pops_flow_mag(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor ii, torch::Tensor jj,
                            torch::Tensor kk, double beta) {
  c10::cuda::CUDAGuard guard(poses.device());
  PopsIn x = pops_inputs(poses, patches, intrinsics, ii, jj, kk);
  torch::Tensor out = torch::empty({1, x.E, x.P, x.P}, x.po.options());
  check_rc(wv_pops_flow_mag(x.po.data_ptr<float>(), x.pa.data_ptr<float>(), x.in.data_ptr<float>(), x.ii.data_ptr<int64_t>(),
                            x.jj.data_ptr<int64_t>(), x.kk.data_ptr<int64_t>(), x.E, x.n_poses, x.n_patches, x.P, (float)beta,
                            out.data_ptr<float>(), status_word(x.po.device()).data_ptr<int>(), current_stream()),
           "cuda_ba.flow_mag");
  if (env_flag("WVB200_CHECK_INDICES")) raise_on_status(x.po.device(), "cuda_ba.flow_mag");
  return out;
}

torch::Tensor pops_point_cloud(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, torch::Tensor ix) {
  check_cuda(poses, "poses"); check_cuda(patches, "patches"); check_cuda(intrinsics, "intrinsics");
  c10::cuda::CUDAGuard guard(poses.device());
  torch::Tensor po = poses.to(torch::kFloat).contiguous(), pa = patches.to(torch::kFloat).contiguous(),
                in = intrinsics.to(torch::kFloat).contiguous(), ixc = idx(ix, "ix");
  const int P = pa.size(-1);
  const int64_t n = pa.numel() / (3 * P * P), n_poses = po.numel() / 7;
  TORCH_CHECK(ixc.numel() == n, "point_cloud: ix must have one frame index per patch");
  torch::Tensor out = torch::empty({1, n, P, P, 4}, po.options());
  check_rc(wv_pops_point_cloud(po.data_ptr<float>(), pa.data_ptr<float>(), in.data_ptr<float>(), ixc.data_ptr<int64_t>(), n, n_poses, P,
                               out.data_ptr<float>(), status_word(po.device()).data_ptr<int>(), current_stream()),
           "cuda_ba.point_cloud");
  if (env_flag("WVB200_CHECK_INDICES")) raise_on_status(po.device(), "cuda_ba.point_cloud");
  return out;
}

torch::Tensor frame_distance(torch::Tensor poses, torch::Tensor patches, torch::Tensor intrinsics, int64_t n, int64_t M, double beta) {
  check_cuda(poses, "poses"); check_cuda(patches, "patches"); check_cuda(intrinsics, "intrinsics");
  c10::cuda::CUDAGuard guard(poses.device());
  torch::Tensor po = poses.to(torch::kFloat).contiguous(), pa = patches.to(torch::kFloat).contiguous(),
                in = intrinsics.to(torch::kFloat).contiguous();
  const int P = pa.size(-1);
  torch::Tensor D = torch::empty({n, n}, po.options());
  check_rc(wv_frame_distance(po.data_ptr<float>(), pa.data_ptr<float>(), in.data_ptr<float>(), n, po.numel() / 7,
                             pa.numel() / (3 * P * P), (int)M, P, (float)beta, D.data_ptr<float>(), current_stream()),
           "cuda_ba.frame_distance");
  return D;
}

std::vector<torch::Tensor> solve_system(torch::Tensor J_Ginv_i, torch::Tensor J_Ginv_j, torch::Tensor ii, torch::Tensor jj,
                                        torch::Tensor res, float ep, float lm, int freen) {
  const torch::Device device = res.device();
  torch::Tensor Ji = J_Ginv_i.to(torch::kCPU, torch::kFloat).contiguous(), Jj = J_Ginv_j.to(torch::kCPU, torch::kFloat).contiguous();
  torch::Tensor ic = ii.to(torch::kCPU, torch::kLong).contiguous(), jc = jj.to(torch::kCPU, torch::kLong).contiguous();
  torch::Tensor rs = res.to(torch::kCPU, torch::kFloat).contiguous();
  const int64_t r = rs.size(0);
  int64_t n = 0;
  int rc = wv_solve_system_host(Ji.data_ptr<float>(), Jj.data_ptr<float>(), ic.data_ptr<int64_t>(), jc.data_ptr<int64_t>(),
                                rs.data_ptr<float>(), r, ep, lm, freen, nullptr, &n);
  TORCH_CHECK(rc == WV_OK, "solve_system: an edge connects a pose to itself or has a negative index");
  torch::Tensor delta = torch::zeros({n, 7}, torch::dtype(torch::kFloat));
  rc = wv_solve_system_host(Ji.data_ptr<float>(), Jj.data_ptr<float>(), ic.data_ptr<int64_t>(), jc.data_ptr<int64_t>(),
                            rs.data_ptr<float>(), r, ep, lm, freen, delta.data_ptr<float>(), &n);
  TORCH_CHECK(rc == WV_OK, "solve_system: ", wv_strerror(rc));
  return {delta.to(device)};
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("forward", &ba_forward, "BA forward operator", py::arg("poses"), py::arg("patches"), py::arg("intrinsics"),
        py::arg("target"), py::arg("weight"), py::arg("lmbda"), py::arg("ii"), py::arg("jj"), py::arg("kk"), py::arg("t0"),
        py::arg("t1"), py::arg("iterations"), py::arg("patches_est") = py::none(), py::arg("mu") = 2.0);
  m.def("global_build", &global_build, "global BA: this shard's partial [S | y]", py::arg("poses"), py::arg("patches"),
        py::arg("intrinsics"), py::arg("target"), py::arg("weight"), py::arg("lmbda"), py::arg("ii"), py::arg("jj"), py::arg("kk"),
        py::arg("t0"), py::arg("t1"), py::arg("shard") = 0, py::arg("nshards") = 1, py::arg("build_plan") = true,
        py::arg("patches_est") = py::none(), py::arg("mu") = 2.0);
  m.def("global_span", &global_span, "global BA: widest frame span of a patch's observations among the free poses (builds the plan)",
        py::arg("poses"), py::arg("patches"), py::arg("ii"), py::arg("jj"), py::arg("kk"), py::arg("t0"), py::arg("t1"), py::arg("build_plan") = true);
  m.def("band_block", &band_block, "half-width (frames) of the band path for a span, 0 = dense path");
  m.def("global_build_band", &global_build_band, "global BA: this shard's partial band buffer [band | y]", py::arg("poses"),
        py::arg("patches"), py::arg("intrinsics"), py::arg("target"), py::arg("weight"), py::arg("lmbda"), py::arg("ii"), py::arg("jj"),
        py::arg("kk"), py::arg("t0"), py::arg("t1"), py::arg("shard"), py::arg("nshards"), py::arg("bs"),
        py::arg("patches_est") = py::none(), py::arg("mu") = 2.0);
  m.def("global_solve_band", &global_solve_band, "global BA: damping + band Cholesky solve -> dX");
  m.def("global_solve", &global_solve, "global BA: damping + dense Cholesky solve -> dX");
  m.def("global_dz", &global_dz, "global BA: this shard's depth updates", py::arg("poses"), py::arg("patches"), py::arg("intrinsics"),
        py::arg("target"), py::arg("weight"), py::arg("lmbda"), py::arg("ii"), py::arg("jj"), py::arg("kk"), py::arg("t0"),
        py::arg("t1"), py::arg("shard"), py::arg("nshards"), py::arg("dX"), py::arg("patches_est") = py::none(), py::arg("mu") = 2.0);
  m.def("global_retract", &global_retract, "global BA: pose + patch retraction");
  m.def("check_status", [](torch::Tensor like) { c10::cuda::CUDAGuard guard(like.device()); raise_on_status(like.device(), "cuda_ba"); },
        "raise if an earlier call on this device flagged a data-dependent failure (bad index, non-SPD system, ...)");
  m.def("neighbors", &neighbors, "temporal neighboor indicies");
  m.def("reproject", &reproject, "fused reprojection of all patch pixels");
  m.def("solve_system", &solve_system, "solve system");
  m.def("transform", &pops_transform, "fused projective transform (projective_ops.transform, jacobian=False)", py::arg("poses"),
        py::arg("patches"), py::arg("intrinsics"), py::arg("ii"), py::arg("jj"), py::arg("kk"), py::arg("depth") = false,
        py::arg("valid") = false, py::arg("tonly") = false);
  m.def("transform_jacobian", &pops_transform_jacobian, "fused projective_ops.transform(jacobian=True): [x1, valid, Ji, Jj, Jz]",
        py::arg("poses"), py::arg("patches"), py::arg("intrinsics"), py::arg("ii"), py::arg("jj"), py::arg("kk"), py::arg("depth") = false,
        py::arg("tonly") = false);
  m.def("flow_mag", &pops_flow_mag, "fused projective_ops.flow_mag", py::arg("poses"), py::arg("patches"), py::arg("intrinsics"),
        py::arg("ii"), py::arg("jj"), py::arg("kk"), py::arg("beta") = 0.3);
  m.def("point_cloud", &pops_point_cloud, "fused projective_ops.point_cloud");
  m.def("frame_distance", &frame_distance, "all-pairs keyframe distance (DPVO.compute_keyframe_distance)", py::arg("poses"),
        py::arg("patches"), py::arg("intrinsics"), py::arg("n"), py::arg("M"), py::arg("beta") = 0.5);
}
