"""GPU (B200) parity tests: the CUDA path, called through the reference-named torch extensions (which go
through the C ABI of include/wvb200.h), against (1) golden outputs of the reference itself and (2) the CPU oracle."""
import ctypes
import importlib
import os

import numpy as np
import pytest
import torch

import oracle
from helpers import PKG, dev, pkg, quat_angle, run_ba, run_corr, vp

pytestmark = pytest.mark.gpu


# ------------------------------------------------------------------------------------------------- corr
@pytest.mark.parametrize("name,radius", [("f32_r3", 3), ("f32_r1", 1)])
def test_corr_fp32_matches_reference(golden, name, radius):
    g = golden["corr_ref"]
    out = run_corr(g[name + "_fmap1"], g[name + "_fmap2"], g[name + "_coords"], g[name + "_ii"], g[name + "_jj"], radius, torch.float32)
    ref = g[name + "_out"]
    # tolerance (stated): fp32 maps, rtol 1e-5 / atol 1e-5 (FMA contraction and summation order only)
    np.testing.assert_allclose(out, ref, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("name,radius", [("f16_r3", 3), ("f16_r0", 0)])
def test_corr_fp16_matches_reference(golden, name, radius):
    g = golden["corr_ref"]
    args = (g[name + "_fmap1"], g[name + "_fmap2"], g[name + "_coords"], g[name + "_ii"], g[name + "_jj"], radius)
    out = run_corr(*args, torch.float16)
    ref = g[name + "_out"]
    truth = oracle.corr_forward(*args, mode=oracle.F64)
    # tolerance (stated): |new - ref| <= 2e-2 * max(1, |ref|): the reference accumulates 128 products in fp16
    assert np.abs(out - ref).max() <= 2e-2 * max(1.0, np.abs(ref).max())
    # and the new path (fp32 accumulate, one final rounding) must be at least as close to the fp64 truth as the reference
    assert np.abs(out - truth).max() <= max(np.abs(ref - truth).max(), 1e-3 * max(1.0, np.abs(truth).max()))
    # one fp16 rounding of the truth: half an ulp of the result
    assert np.abs(out - truth).max() <= 2.0 ** -10 * max(1.0, np.abs(truth).max())


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-6), (torch.float16, 2.0 ** -10)])
def test_corr_tracker_workload_vs_oracle(syn, dtype, tol):
    w = syn.make_corr_workload(seed=1, M=8, pmem=12, n_frames=14, ht=24, wd=32, C=128, lifetime=6, removal=9, oob_frac=0.1)
    for lvl, (fm, div) in enumerate(((w["fmap0"], 1.0), (w["fmap1"], 4.0))):
        g = torch.from_numpy(w["gmap"]).to(dtype).float().numpy()
        f = torch.from_numpy(fm).to(dtype).float().numpy()
        co = (w["coords"] / div).astype(np.float32)
        out = run_corr(g, f, co, w["ii1"], w["jj1"], 3, dtype)
        truth = oracle.corr_forward(g, f, co, w["ii1"], w["jj1"], 3, mode=oracle.F64)
        assert np.abs(out - truth).max() <= tol * max(1.0, np.abs(truth).max()), f"level {lvl}"


def test_corr_empty_and_bad_index(golden):
    cc, _ = pkg().native_modules()
    g = golden["corr_ref"]
    f1, f2 = dev(g["f32_r3_fmap1"]), dev(g["f32_r3_fmap2"])
    out, = cc.forward(f1, f2, torch.zeros(1, 0, 2, 3, 3, device="cuda"), torch.zeros(0, dtype=torch.long, device="cuda"),
                      torch.zeros(0, dtype=torch.long, device="cuda"), 3)
    assert tuple(out.shape) == (1, 0, 7, 7, 3, 3)
    os.environ["WVB200_CHECK_INDICES"] = "1"
    try:
        with pytest.raises(IndexError):
            cc.forward(f1, f2, dev(g["f32_r3_coords"][:, :2]), torch.tensor([0, 99], device="cuda"), torch.tensor([0, 0], device="cuda"), 3)
    finally:
        del os.environ["WVB200_CHECK_INDICES"]


@pytest.mark.parametrize("name,radius", [("f32_r3", 3), ("f32_r1", 1), ("f16_r3", 3)])
def test_corr_backward_matches_reference(golden, name, radius):
    cc, _ = pkg().native_modules()
    g = golden["corr_ref"]
    g1, g2 = cc.backward(dev(g[name + "_fmap1"]), dev(g[name + "_fmap2"]), dev(g[name + "_coords"]), dev(g[name + "_ii"]),
                         dev(g[name + "_jj"]), dev(g[name + "_grad"]), radius)
    for mine, ref in ((g1, g[name + "_g1"]), (g2, g[name + "_g2"])):
        assert np.abs(mine.cpu().numpy() - ref).max() <= 5e-5 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("r", [0, 1, 3])
def test_patchify_matches_reference(golden, r):
    cc, _ = pkg().native_modules()
    g = golden["corr_ref"]
    p, = cc.patchify_forward(dev(g["patchify_net"]), dev(g["patchify_coords"]), r)
    assert np.array_equal(p.cpu().numpy(), g[f"patchify_r{r}_out"])  # copies: bit-exact
    ph, = cc.patchify_forward(dev(g["patchify_net"]).half(), dev(g["patchify_coords"]), r)
    assert np.array_equal(ph.float().cpu().numpy(), torch.from_numpy(g[f"patchify_r{r}_out"]).half().float().numpy())
    ng, = cc.patchify_backward(dev(g["patchify_net"]), dev(g["patchify_coords"]), dev(g[f"patchify_r{r}_grad"]), r)
    assert np.abs(ng.cpu().numpy() - g[f"patchify_r{r}_netgrad"]).max() < 1e-5


def test_altcorr_python_mirror_autograd(golden):
    """the host-side mirror of dpvo/altcorr/correlation.py: corr()/patchify() with autograd"""
    ac = pkg().altcorr
    g = golden["corr_ref"]
    f1 = dev(g["f32_r3_fmap1"]).requires_grad_(True)
    f2 = dev(g["f32_r3_fmap2"]).requires_grad_(True)
    out = ac.corr(f1, f2, dev(g["f32_r3_coords"]), dev(g["f32_r3_ii"]), dev(g["f32_r3_jj"]), 3)
    (out * dev(g["f32_r3_grad"])).sum().backward()
    assert np.abs(f1.grad.cpu().numpy() - g["f32_r3_g1"]).max() <= 5e-5 * np.abs(g["f32_r3_g1"]).max()
    assert np.abs(f2.grad.cpu().numpy() - g["f32_r3_g2"]).max() <= 5e-5 * np.abs(g["f32_r3_g2"]).max()
    net = dev(g["patchify_net"])
    co = dev(g["patchify_coords"])
    pt = ac.patchify(net, co, 1)
    assert tuple(pt.shape) == (2, 7, 5, 3, 3)
    crop = torch.from_numpy(g["patchify_r1_out"]).cuda()
    fr = co - co.floor()
    dx, dy = fr[..., 0][:, :, None, None, None], fr[..., 1][:, :, None, None, None]
    want = (1 - dy) * (1 - dx) * crop[..., :3, :3] + (1 - dy) * dx * crop[..., :3, 1:] + dy * (1 - dx) * crop[..., 1:, :3] + dy * dx * crop[..., 1:, 1:]
    assert torch.allclose(pt, want, atol=1e-5)


# ------------------------------------------------------------------------------------------------- BA
def _golden_scene(g, name):
    return {k: g[f"{name}_in_{k}"] for k in ("poses", "patches", "intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk", "patches_est")}


@pytest.mark.parametrize("name", ["win", "trk", "off"])
@pytest.mark.parametrize("tag", ["full", "late", "struct"])
def test_ba_matches_reference(golden, name, tag):
    g = golden["ba_ref"]
    s = _golden_scene(g, name)
    t0, t1, iters = (int(v) for v in g[f"{name}_{tag}_args"])
    poses, patches = run_ba(s, t0, t1, iters)
    rp, rd = g[f"{name}_{tag}_poses"], g[f"{name}_{tag}_patches"]
    # tolerances (stated): translation atol 2e-4, rotation 1e-4 rad, inverse depth rtol 2e-3 after `iters` iterations
    assert np.abs(poses[:, :3] - rp[:, :3]).max() < 2e-4
    assert quat_angle(poses[:, 3:], rp[:, 3:]).max() < 1e-3
    assert np.abs(patches[:, 2] - rd[:, 2]).max() < 2e-3 * max(1.0, np.abs(rd[:, 2]).max())
    assert np.array_equal(patches[:, :2], rd[:, :2])  # x, y channels untouched
    # every PxP entry of the depth channel carries the same value (patch_retr_kernel)
    assert np.array_equal(patches[:, 2], np.broadcast_to(patches[:, 2, :1, :1], patches[:, 2].shape))


@pytest.mark.parametrize("name", ["lc257", "lc40", "lc1"])
def test_ba_loop_closure_call_shape_matches_reference(golden, name):
    """fastba.BA as dpvo/loop_closure/long_term.py:104-127 calls it (3 fixed poses, t0 = t1 = 3, 6 iterations, lmbda 1e-3, two edges
    per keypoint with kk = arange(n).repeat(2) -- NOT grouped by patch --, ii == 1, jj in {0, 2}, full-resolution intrinsics):
    goldens from the reference's own ba_cuda.cu (oracle/make_golden_loop.py).  Structure only: poses must come back bit for bit."""
    g = golden["ba_loop_ref"]
    s = {k: g[f"{name}_in_{k}"] for k in ("poses", "patches", "intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")}
    poses, patches = run_ba(s, 3, 3, 6)
    assert np.array_equal(poses, s["poses"])
    rd = g[f"{name}_patches"]
    # six Gauss-Newton steps on a 1-D problem per point: fp32 round-off only (stated: 1e-4 relative inverse depth)
    assert np.abs(patches[:, 2] - rd[:, 2]).max() < 1e-4 * max(1.0, np.abs(rd[:, 2]).max())
    assert np.array_equal(patches[:, :2], rd[:, :2])
    assert np.array_equal(patches[:, 2], np.broadcast_to(patches[:, 2, :1, :1], patches[:, 2].shape))
    poses2, patches2 = run_ba(s, 3, 3, 6)
    assert np.array_equal(patches, patches2)                              # deterministic


@pytest.mark.parametrize("M,edges", [(96, "tracker"), (32, "all")])
def test_ba_full_size_vs_oracle_and_deterministic(syn, M, edges):
    n = 40 if edges == "tracker" else 10
    s = syn.make_ba_scene(seed=2, n=n, M=M, edges=edges)
    t1 = s["t1"]
    t0 = t1 - 10 if edges == "tracker" else s["t0"]
    o = oracle.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"],
                          s["kk"], t0, t1, 2, mode=oracle.F64)
    assert o["status"] == 0
    # the same arithmetic in fp32 in the reference's own evaluation order: how far fp32 rounding alone moves the
    # answer on this scene (the all-pairs window is ill conditioned: 4e-4; the tracker graph: 4e-6)
    o32 = oracle.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"],
                            s["kk"], t0, t1, 2, mode=oracle.F32)
    poses, patches = run_ba(s, t0, t1, 2)
    # tolerance (stated): translation max(1e-4, the reference-order fp32 deviation from fp64), rotation likewise in rad,
    # inverse depth rtol 1e-3
    tol_t = max(1e-4, np.abs(o32["poses"][:, :3] - o["poses"][:, :3]).max())
    tol_r = max(1e-4, quat_angle(o32["poses"][:, 3:], o["poses"][:, 3:]).max())
    assert np.abs(poses[:, :3] - o["poses"][:, :3]).max() < tol_t
    assert quat_angle(poses[:, 3:], o["poses"][:, 3:]).max() < tol_r
    d, do = patches[:, 2, 0, 0], o["patches"][:, 2, 0, 0]
    assert np.abs(d - do).max() < 1e-3 * max(1.0, np.abs(do).max())
    poses2, patches2 = run_ba(s, t0, t1, 2)
    assert np.array_equal(poses, poses2) and np.array_equal(patches, patches2)  # fixed summation order
    os.environ["WVB200_NO_COOP"] = "1"  # multi-launch plan must give the same bits
    try:
        poses3, patches3 = run_ba(s, t0, t1, 2)
    finally:
        del os.environ["WVB200_NO_COOP"]
    assert np.array_equal(poses, poses3) and np.array_equal(patches, patches3)


def test_ba_general_graph_slow_path(syn):
    """duplicate (patch, frame) edges and patches observed from two different source frames: the serialised path"""
    s = syn.make_ba_scene(seed=4, n=7, M=5, edges="all")
    rng = np.random.default_rng(0)
    dup = rng.integers(0, len(s["kk"]), 40)
    for k in ("ii", "jj", "kk", "target", "weight"):
        s[k] = np.concatenate([s[k], s[k][dup]])
    s["ii"][-10:] = (s["ii"][-10:] + 1) % 7  # inconsistent source frame for some patches
    perm = rng.permutation(len(s["kk"]))
    for k in ("ii", "jj", "kk", "target", "weight"):
        s[k] = np.ascontiguousarray(s[k][perm])
    o = oracle.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"],
                          s["kk"], s["t0"], s["t1"], 2, mode=oracle.F64)
    poses, patches = run_ba(s, s["t0"], s["t1"], 2)
    assert np.abs(poses - o["poses"]).max() < 2e-4
    assert np.abs(patches[:, 2, 0, 0] - o["patches"][:, 2, 0, 0]).max() < 2e-3


def test_ba_many_edges_per_patch(syn):
    """segments longer than a warp (> 32 edges per patch)"""
    s = syn.make_ba_scene(seed=6, n=40, M=3, edges="all", pose_noise=0.005)
    t0, t1 = s["t1"] - 8, s["t1"]
    o = oracle.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"],
                          s["kk"], t0, t1, 2, mode=oracle.F64)
    poses, patches = run_ba(s, t0, t1, 2)
    assert np.abs(poses - o["poses"]).max() < 2e-4
    assert np.abs(patches[:, 2, 0, 0] - o["patches"][:, 2, 0, 0]).max() < 2e-3


@pytest.mark.parametrize("name", ["win", "trk"])
def test_ba_depth_prior_term(golden, syn, name):
    kw = {"win": dict(seed=0, n=6, M=8, edges="all"), "trk": dict(seed=3, n=12, M=6, edges="tracker", lifetime=5, removal=8)}[name]
    s = syn.make_ba_scene(**kw)
    poses, patches = run_ba(s, s["t0"], s["t1"], 1, patches_est=s["patches_est"], mu=2.0)
    o = oracle.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"],
                          s["kk"], s["t0"], s["t1"], 1, patches_est=s["patches_est"], mu=2.0, mode=oracle.F64)
    assert np.abs(poses - o["poses"]).max() < 1e-4
    assert np.abs(patches[:, 2, 0, 0] - o["patches"][:, 2, 0, 0]).max() < 1e-3
    rp = golden["pyba_ref"][f"{name}_prior_poses"]  # the reference's own dpvo/ba.py
    assert np.abs(poses[:, :3] - rp[:, :3]).max() < 3e-4


def test_ba_not_spd_raises_and_recovers(syn):
    _, cb = pkg().native_modules()
    s = syn.make_ba_scene(seed=0, n=6, M=8, edges="all")
    bad = dict(s); bad["weight"] = np.full_like(s["weight"], np.nan)
    with pytest.raises(RuntimeError):
        run_ba(bad, s["t0"], s["t1"], 2)
    poses, _ = run_ba(s, s["t0"], s["t1"], 2)  # status word was cleared: next call is healthy
    assert np.isfinite(poses).all()
    with pytest.raises(IndexError):
        worse = dict(s); worse["kk"] = s["kk"].copy(); worse["kk"][3] = 10 ** 7
        run_ba(worse, s["t0"], s["t1"], 2)


@pytest.mark.parametrize("name", ["win", "trk", "off"])
def test_reproject_and_neighbors_match_reference(golden, name):
    _, cb = pkg().native_modules()
    g = golden["ba_ref"]
    s = _golden_scene(g, name)
    rp = cb.reproject(dev(s["poses"]), dev(s["patches"]), dev(s["intrinsics"]), dev(s["ii"]), dev(s["jj"]), dev(s["kk"]))
    assert tuple(rp.shape) == g[f"{name}_reproject"].shape
    np.testing.assert_allclose(rp.cpu().numpy(), g[f"{name}_reproject"], rtol=1e-4, atol=2e-3)
    ix, jx = cb.neighbors(dev(s["kk"]), dev(s["jj"]))
    assert ix.dtype == torch.int64 and ix.is_cuda
    assert np.array_equal(ix.cpu().numpy(), g[f"{name}_nb_ix"]) and np.array_equal(jx.cpu().numpy(), g[f"{name}_nb_jx"])


def test_neighbors_bit_exact_random_and_full_size(golden, syn):
    _, cb = pkg().native_modules()
    g = golden["ba_ref"]
    ix, jx = cb.neighbors(dev(g["nb_rand_ii"]), dev(g["nb_rand_jj"]))
    assert np.array_equal(ix.cpu().numpy(), g["nb_rand_ix"]) and np.array_equal(jx.cpu().numpy(), g["nb_rand_jx"])
    ii, jj, kk = syn.tracker_edges(40, 96)
    rng = np.random.default_rng(5)
    perm = rng.permutation(len(kk))
    for a, b in ((kk, jj), (kk[perm], jj[perm]), (kk + 700000, jj)):
        ix, jx = cb.neighbors(dev(a), dev(b))
        oi, oj = oracle.neighbors(a, b)
        assert np.array_equal(ix.cpu().numpy(), oi) and np.array_equal(jx.cpu().numpy(), oj)


def test_solve_system_matches_reference(golden):
    _, cb = pkg().native_modules()
    g = golden["ba_ref"]
    for tag, freen in (("full", -1), ("free4", 4)):
        d, = cb.solve_system(dev(g["pgo_Ji"]), dev(g["pgo_Jj"]), dev(g["pgo_ii"]), dev(g["pgo_jj"]), dev(g["pgo_res"]), 1e-4, 1e-4, freen)
        assert d.is_cuda
        assert np.abs(d.cpu().numpy() - g[f"pgo_{tag}_delta"]).max() < 1e-5


# ------------------------------------------------------------------------------------------------- C ABI, host buffers
def test_c_abi_host_buffer_entries(golden, syn):
    lib = pkg().capi()
    assert lib.wv_device_supported() == 1
    g = golden["corr_ref"]
    f1, f2, co = (np.ascontiguousarray(g["f32_r3_" + k]) for k in ("fmap1", "fmap2", "coords"))
    ii, jj = np.ascontiguousarray(g["f32_r3_ii"]), np.ascontiguousarray(g["f32_r3_jj"])
    out = np.empty_like(g["f32_r3_out"])
    rc = lib.wv_corr_forward_host(vp(f1), vp(f2), vp(co), vp(ii), vp(jj), vp(out), 1, 1, co.shape[1], f1.shape[2], 3, 3,
                                  f1.shape[1], f2.shape[1], f2.shape[3], f2.shape[4], 3, None)
    assert rc == 0
    np.testing.assert_allclose(out, g["f32_r3_out"], rtol=1e-5, atol=1e-5)
    s = syn.make_ba_scene(seed=0, n=6, M=8, edges="all")
    poses, patches = s["poses"].copy(), s["patches"].copy()
    st = ctypes.c_int(0)
    f = lib.wv_ba_forward_host
    f.argtypes = [ctypes.c_void_p] * 9 + [ctypes.c_int64] * 3 + [ctypes.c_int] * 4 + [ctypes.c_void_p, ctypes.c_float,
                                                                                      ctypes.c_void_p, ctypes.c_void_p]
    rc = f(vp(poses), vp(patches), vp(s["intrinsics"]), vp(s["target"]), vp(s["weight"]), vp(s["lmbda"]), vp(s["ii"]), vp(s["jj"]),
           vp(s["kk"]), len(s["kk"]), poses.shape[0], patches.shape[0], 3, s["t0"], s["t1"], 2, None, 2.0, ctypes.byref(st), None)
    assert rc == 0 and st.value == 0
    o = oracle.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"],
                          s["kk"], s["t0"], s["t1"], 2, mode=oracle.F64)
    assert np.abs(poses - o["poses"]).max() < 1e-4


# ------------------------------------------------------------------------------------------------- tiled (TMA + mma) corr
def _tiled(on):
    class _Ctx:
        def __enter__(self):
            self.old = os.environ.get("WVB200_CORR_TILED")
            os.environ["WVB200_CORR_TILED"] = "1" if on else "0"

        def __exit__(self, *a):
            if self.old is None:
                del os.environ["WVB200_CORR_TILED"]
            else:
                os.environ["WVB200_CORR_TILED"] = self.old
    return _Ctx()


def _check_tiled(g, f, co, ii, jj):
    """tiled path vs fp64 oracle (half an fp16 ulp of the result scale) and vs the generic kernel"""
    with _tiled(True):
        out = run_corr(g, f, co, ii, jj, 3, torch.float16)
    with _tiled(False):
        gen = run_corr(g, f, co, ii, jj, 3, torch.float16)
    os.environ["WVB200_CORR_TC5"] = "0"      # the mma.sync form of the tile kernel: same tiles, same blend, other tensor path
    try:
        with _tiled(True):
            legacy = run_corr(g, f, co, ii, jj, 3, torch.float16)
    finally:
        del os.environ["WVB200_CORR_TC5"]
    g16 = torch.from_numpy(g).half().float().numpy()
    f16 = torch.from_numpy(f).half().float().numpy()
    truth = oracle.corr_forward(g16, f16, co, ii, jj, 3, mode=oracle.F64)
    scale = max(1.0, np.abs(truth).max())
    assert np.abs(out - truth).max() <= 2.0 ** -10 * scale
    # both paths accumulate in fp32 and round once: they may differ by one fp16 rounding where the sums differ in the last bits
    assert np.abs(out - gen).max() <= 2.0 ** -9 * scale
    assert np.abs(legacy - truth).max() <= 2.0 ** -10 * scale
    assert np.abs(out - legacy).max() <= 2.0 ** -9 * scale
    return out


@pytest.mark.parametrize("lvl", [0, 1])
def test_corr_tiled_tracker_workload(syn, lvl):
    w = syn.make_corr_workload(seed=1, M=8, pmem=12, n_frames=14, ht=24, wd=32, C=128, lifetime=6, removal=9, oob_frac=0.15)
    fm, div = ((w["fmap0"], 1.0), (w["fmap1"], 4.0))[lvl]
    _check_tiled(w["gmap"], fm, (w["coords"] / div).astype(np.float32), w["ii1"], w["jj1"])


def test_corr_tiled_magnified_patches_and_odd_map(syn):
    """patch spacing 1.0 .. 3.5 px: small box, large box and the direct-gather path; 23x31 map (odd HW: scalar restaging)"""
    rng = np.random.default_rng(7)
    n1, n2, H, W, E = 40, 5, 23, 31, 600
    g = (0.25 * rng.standard_normal((1, n1, 128, 3, 3))).astype(np.float32)
    f = (0.25 * rng.standard_normal((1, n2, 128, H, W))).astype(np.float32)
    ctr = np.stack([rng.uniform(-6, W + 6, E), rng.uniform(-6, H + 6, E)], -1)
    sp = rng.uniform(1.0, 3.5, E)
    gy, gx = np.meshgrid(np.arange(-1, 2), np.arange(-1, 2), indexing="ij")
    co = np.stack([ctr[:, 0, None, None] + sp[:, None, None] * gx, ctr[:, 1, None, None] + sp[:, None, None] * gy], 1)[None].astype(np.float32)
    ii = rng.integers(0, n1, E); jj = rng.integers(0, n2, E)
    ii[100:130] = ii[100]  # runs of edges that share a patch (A-fragment reuse)
    out = _check_tiled(g, f, co, ii.astype(np.int64), jj.astype(np.int64))
    assert np.abs(out).max() > 0.1


def test_corr_tiled_every_window_needs_two_slots(syn):
    """spacing 1.6 px: every edge is a 12x12 box in two half tiles, so a producer batch of 16 edges wants 32 record slots --
    more than the record ring may hand out at once (the batch is cut; see the producer)"""
    rng = np.random.default_rng(21)
    n1, n2, H, W, E = 30, 4, 48, 64, 1500
    g = (0.25 * rng.standard_normal((1, n1, 128, 3, 3))).astype(np.float32)
    f = (0.25 * rng.standard_normal((1, n2, 128, H, W))).astype(np.float32)
    f1 = (0.25 * rng.standard_normal((1, n2, 128, H // 4, W // 4))).astype(np.float32)
    ctr = np.stack([rng.uniform(4, W - 4, E), rng.uniform(4, H - 4, E)], -1)
    gy, gx = np.meshgrid(np.arange(-1, 2), np.arange(-1, 2), indexing="ij")
    co = np.stack([ctr[:, 0, None, None] + 1.6 * gx, ctr[:, 1, None, None] + 1.6 * gy], 1)[None].astype(np.float32)
    ii = rng.integers(0, n1, E).astype(np.int64); jj = rng.integers(0, n2, E).astype(np.int64)
    _check_tiled(g, f, co, ii, jj)
    _check_stacked(g, f, f1, co, ii, jj)


def test_corr_tiled_batch_and_few_edges(syn):
    """B = 2 and fewer edges than consumer warps"""
    rng = np.random.default_rng(8)
    n1, n2, H, W, E = 6, 3, 24, 32, 5
    g = (0.25 * rng.standard_normal((2, n1, 128, 3, 3))).astype(np.float32)
    f = (0.25 * rng.standard_normal((2, n2, 128, H, W))).astype(np.float32)
    c = rng.uniform(2, 20, (2, E, 2))
    gy, gx = np.meshgrid(np.arange(-1, 2), np.arange(-1, 2), indexing="ij")
    co = np.stack([c[..., 0, None, None] + gx, c[..., 1, None, None] + gy], 2).astype(np.float32)
    _check_tiled(g, f, co, rng.integers(0, n1, E).astype(np.int64), rng.integers(0, n2, E).astype(np.int64))


def _check_stacked(g, f0, f1, co, ii, jj):
    """two-level kernel vs the two-call form (same tiles, same arithmetic: bit-identical) and vs the fp64 oracle"""
    cc, _ = pkg().native_modules()
    altcorr = importlib.import_module(PKG + ".altcorr.correlation")
    gd, f0d, f1d, cod, iid, jjd = dev(g, torch.float16), dev(f0, torch.float16), dev(f1, torch.float16), dev(co), dev(ii), dev(jj)
    st = altcorr.corr_stacked(gd, (f0d, f1d), cod, iid, jjd, 3)
    B, E = co.shape[0], co.shape[1]
    assert st.shape == (B, E, 882) and st.dtype == torch.float16
    with _tiled(True):
        c0, = cc.forward(gd, f0d, cod / 1, iid, jjd, 3)
        c1, = cc.forward(gd, f1d, cod / 4, iid, jjd, 3)
    two = torch.stack([c0, c1], -1).view(B, E, -1)   # dpvo/dpvo.py:333
    torch.cuda.synchronize()
    assert torch.equal(st, two)
    g16, f016, f116 = (torch.from_numpy(x).half().float().numpy() for x in (g, f0, f1))
    t0 = oracle.corr_forward(g16, f016, co, ii, jj, 3, mode=oracle.F64)
    t1 = oracle.corr_forward(g16, f116, (co / 4).astype(np.float32), ii, jj, 3, mode=oracle.F64)
    truth = np.stack([t0, t1], -1).reshape(B, E, -1)
    scale = max(1.0, np.abs(truth).max())
    assert np.abs(st.float().cpu().numpy() - truth).max() <= 2.0 ** -10 * scale
    return st


def test_corr_stacked_tracker_workload(syn):
    w = syn.make_corr_workload(seed=2, M=8, pmem=12, n_frames=14, ht=24, wd=32, C=128, lifetime=6, removal=9, oob_frac=0.15)
    st = _check_stacked(w["gmap"], w["fmap0"], w["fmap1"], w["coords"], w["ii1"], w["jj1"])
    assert st.float().abs().max() > 0.1


def test_corr_stacked_magnified_batch_and_few_edges(syn):
    """small / large / direct-gather windows at both levels, B = 2, odd maps, fewer edges than pipelines"""
    rng = np.random.default_rng(17)
    n1, n2, E = 40, 5, 500
    gy, gx = np.meshgrid(np.arange(-1, 2), np.arange(-1, 2), indexing="ij")
    for B, H0, W0, H1, W1, E in ((1, 46, 62, 11, 15, 500), (2, 24, 32, 6, 8, 3)):
        g = (0.25 * rng.standard_normal((B, n1, 128, 3, 3))).astype(np.float32)
        f0 = (0.25 * rng.standard_normal((B, n2, 128, H0, W0))).astype(np.float32)
        f1 = (0.25 * rng.standard_normal((B, n2, 128, H1, W1))).astype(np.float32)
        ctr = np.stack([rng.uniform(-8, W0 + 8, (B, E)), rng.uniform(-8, H0 + 8, (B, E))], -1)
        sp = rng.uniform(1.0, 14.0, (B, E))        # up to 3.5 px spacing at level 1, far beyond the tiles at level 0
        co = np.stack([ctr[..., 0, None, None] + sp[..., None, None] * gx, ctr[..., 1, None, None] + sp[..., None, None] * gy], 2).astype(np.float32)
        ii = rng.integers(0, n1, E).astype(np.int64); jj = rng.integers(0, n2, E).astype(np.int64)
        if E > 200:
            ii[100:130] = ii[100]
        _check_stacked(g, f0, f1, co, ii, jj)


def test_corr_stacked_composed_form_and_bad_index(golden):
    """fp32 maps take the composed two-call form (same result layout); a bad index raises from the fused call"""
    cc, _ = pkg().native_modules()
    altcorr = importlib.import_module(PKG + ".altcorr.correlation")
    rng = np.random.default_rng(19)
    g = dev((0.25 * rng.standard_normal((1, 4, 128, 3, 3))).astype(np.float32))
    f0 = dev((0.25 * rng.standard_normal((1, 2, 128, 24, 32))).astype(np.float32))
    f1 = dev((0.25 * rng.standard_normal((1, 2, 128, 6, 8))).astype(np.float32))
    co = dev(np.full((1, 3, 2, 3, 3), 10.25, np.float32))
    ii, jj = torch.tensor([0, 1, 2], device="cuda"), torch.tensor([0, 1, 1], device="cuda")
    comp = altcorr.corr_stacked(g, (f0, f1), co, ii, jj, 3)
    fused = altcorr.corr_stacked(g.half(), (f0.half(), f1.half()), co, ii, jj, 3)
    assert comp.shape == fused.shape == (1, 3, 882)
    assert (comp - fused.float()).abs().max() < 2e-2
    empty = altcorr.corr_stacked(g.half(), (f0.half(), f1.half()), co[:, :0], ii[:0], jj[:0], 3)
    assert empty.shape == (1, 0, 882)
    # non-finite and far-out-of-map coordinates: zeros for that edge like the two-call form, the other edges untouched
    cb = co.clone(); cb[0, 1] = float("inf"); cb[0, 2, 0] = 1e9
    with _tiled(True):
        two = torch.stack([cc.forward(g.half(), f0.half(), cb / 1, ii, jj, 3)[0], cc.forward(g.half(), f1.half(), cb / 4, ii, jj, 3)[0]], -1).view(1, 3, -1)
    st = altcorr.corr_stacked(g.half(), (f0.half(), f1.half()), cb, ii, jj, 3)
    assert torch.isfinite(st[0, 0]).all() and torch.equal(st[0, 0], fused[0, 0]) and torch.equal(torch.nan_to_num(st), torch.nan_to_num(two))
    os.environ["WVB200_CHECK_INDICES"] = "1"
    try:
        with pytest.raises(IndexError):
            cc.forward_stacked(g.half(), f0.half(), f1.half(), co, torch.tensor([0, 9, 2], device="cuda"), jj, 3, 1.0, 4.0)
    finally:
        del os.environ["WVB200_CHECK_INDICES"]
    with pytest.raises(RuntimeError):
        cc.forward_stacked(g.half(), f0.half(), f1.half(), co, ii, jj, 3, 1.0, 3.0)


def test_corr_tiled_bad_index_and_nonfinite(golden):
    cc, _ = pkg().native_modules()
    rng = np.random.default_rng(9)
    g = dev((0.25 * rng.standard_normal((1, 4, 128, 3, 3))).astype(np.float32), torch.float16)
    f = dev((0.25 * rng.standard_normal((1, 2, 128, 24, 32))).astype(np.float32), torch.float16)
    co = np.full((1, 3, 2, 3, 3), 10.25, np.float32)
    co[0, 1] = np.inf
    os.environ["WVB200_CHECK_INDICES"] = "1"
    try:
        with _tiled(True):
            out, = cc.forward(g, f, dev(co), torch.tensor([0, 1, 2], device="cuda"), torch.tensor([0, 1, 1], device="cuda"), 3)
            torch.cuda.synchronize()
            assert torch.isfinite(out[0, 0]).all() and torch.isfinite(out[0, 2]).all()
            with pytest.raises(IndexError):
                cc.forward(g, f, dev(co), torch.tensor([0, 9, 2], device="cuda"), torch.tensor([0, 1, 1], device="cuda"), 3)
    finally:
        del os.environ["WVB200_CHECK_INDICES"]


def test_host_session_update_matches_device_path(syn):
    """hostapi.Session (wv_session_update_host): host buffers in/out, three streams; same bits as the tensor path"""
    w = syn.make_corr_workload(seed=3, M=8, pmem=12, n_frames=14, ht=24, wd=32, C=128, lifetime=6, removal=9, oob_frac=0.1)
    s = syn.make_ba_scene(seed=2, n=14, M=8, edges="tracker", lifetime=6, removal=9)
    E = w["E"]
    assert E == len(s["kk"])
    g16 = torch.from_numpy(w["gmap"]).half(); f0 = torch.from_numpy(w["fmap0"]).half(); f1 = torch.from_numpy(w["fmap1"]).half()
    co = torch.from_numpy(w["coords"]); ii1 = torch.from_numpy(w["ii1"]); jj1 = torch.from_numpy(w["jj1"])
    out0 = torch.empty((1, E, 7, 7, 3, 3), dtype=torch.half); out1 = torch.empty_like(out0)
    T = lambda k: torch.from_numpy(s[k].copy())  # noqa: E731
    t0, t1 = int(s["t1"]) - 6, int(s["t1"])
    ba = dict(poses=T("poses"), patches=T("patches"), intrinsics=T("intrinsics"), target=T("target"), weight=T("weight"),
              lmbda=T("lmbda"), ii=T("ii"), jj=T("jj"), kk=T("kk"), t0=t0, t1=t1,
              nb_ix=torch.empty(E, dtype=torch.long), nb_jx=torch.empty(E, dtype=torch.long))
    sess = pkg().hostapi.Session()
    for forced in ("1", "0"):  # tiled and generic corr
        os.environ["WVB200_CORR_TILED"] = forced
        try:
            ba["poses"].copy_(T("poses")); ba["patches"].copy_(T("patches"))
            sess.update(g16, [f0, f1], co, ii1, jj1, [out0, out1], ba=ba, iterations=2)
            c0 = run_corr(g16.float().numpy(), f0.float().numpy(), w["coords"], w["ii1"], w["jj1"], 3, torch.float16)
            c1 = run_corr(g16.float().numpy(), f1.float().numpy(), (w["coords"] * 0.25).astype(np.float32), w["ii1"], w["jj1"], 3, torch.float16)
        finally:
            del os.environ["WVB200_CORR_TILED"]
        assert np.array_equal(out0.float().numpy(), c0) and np.array_equal(out1.float().numpy(), c1)
        poses, patches = run_ba(s, t0, t1, 2)
        assert np.array_equal(ba["poses"].numpy(), poses) and np.array_equal(ba["patches"].numpy(), patches)
        oi, oj = oracle.neighbors(s["kk"], s["jj"])
        assert np.array_equal(ba["nb_ix"].numpy(), oi) and np.array_equal(ba["nb_jx"].numpy(), oj)
    # the same update with both levels as one stacked volume (DPVO.corr's layout): out0 / out1 above came from the tiled kernel last? no --
    # recompute the tiled per-level outputs and compare the interleaved pairs bit for bit
    os.environ["WVB200_CORR_TILED"] = "1"
    try:
        c0 = run_corr(g16.float().numpy(), f0.float().numpy(), w["coords"], w["ii1"], w["jj1"], 3, torch.float16)
        c1 = run_corr(g16.float().numpy(), f1.float().numpy(), (w["coords"] * 0.25).astype(np.float32), w["ii1"], w["jj1"], 3, torch.float16)
    finally:
        del os.environ["WVB200_CORR_TILED"]
    st = torch.empty((1, E, 882), dtype=torch.half)
    ba["poses"].copy_(T("poses")); ba["patches"].copy_(T("patches"))
    sess.update(g16, [f0, f1], co, ii1, jj1, None, ba=ba, iterations=2, corr_stacked_out=st)
    assert np.array_equal(st.float().numpy(), np.stack([c0, c1], -1).reshape(1, E, 882))
    assert np.array_equal(ba["poses"].numpy(), poses)
    with pytest.raises(RuntimeError):   # fp32 maps: the two-level kernel does not cover them, the session says so
        sess.update(g16.float(), [f0.float(), f1.float()], co, ii1, jj1, None, corr_stacked_out=st.float(), ba=None)
    sess.close()


def test_ba_reuses_the_plan_of_neighbors(syn):
    """neighbors(kk, jj) then BA(..., ii, jj, kk) on the same tensors (one tracker update()): BA runs on the cached plan,
    with the same bits as a BA that builds its own; an in-place edit of kk in between invalidates the cache"""
    _, cb = pkg().native_modules()
    s = syn.make_ba_scene(seed=2, n=40, M=32, edges="tracker")
    t0, t1 = int(s["t1"]) - 10, int(s["t1"])
    D = {k: dev(s[k]) for k in ("intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")}

    def run(cache, edit=False):
        poses, patches = dev(s["poses"].copy()), dev(s["patches"].copy())
        if not cache:
            os.environ["WVB200_NO_PLAN_CACHE"] = "1"
        try:
            ix, jx = cb.neighbors(D["kk"], D["jj"])
            if edit:
                D["kk"].add_(0)  # bumps the version counter: the cached plan must not be trusted
            cb.forward(poses, patches, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, 2)
        finally:
            os.environ.pop("WVB200_NO_PLAN_CACHE", None)
        torch.cuda.synchronize()
        return poses.cpu().numpy(), patches.cpu().numpy(), ix.cpu().numpy()

    a, b, c = run(True), run(False), run(True, edit=True)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    assert np.array_equal(a[0], c[0]) and np.array_equal(a[1], c[1])
    # an out-of-range frame index is caught by the BA kernel itself on the cached-plan path
    bad = D["jj"].clone(); bad[5] = 10 ** 6
    cb.neighbors(D["kk"], bad)
    with pytest.raises(IndexError):
        cb.forward(dev(s["poses"].copy()), dev(s["patches"].copy()), D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], bad,
                   D["kk"], t0, t1, 2)


# ------------------------------------------------------------------------------------------------- global BA (N > 16 free poses)
def _global_tol(s, t0, t1, iters, **kw):
    o = oracle.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"],
                          s["kk"], t0, t1, iters, mode=oracle.F64, **kw)
    o32 = oracle.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"],
                            s["kk"], t0, t1, iters, mode=oracle.F32, **kw)
    assert o["status"] == 0
    # tolerance (stated): max(floor, 3x the deviation of the reference-order fp32 evaluation from the fp64 oracle); floors
    # 2e-4 m / 2e-4 rad / 2e-3 relative inverse depth.  With only pose 0 fixed (t0 == 1, the reference's global BA call) the
    # monocular scale is a gauge direction held only by the +1 damping, so fp32 rounding alone moves the answer by up to ~1e-3
    # (depth ~1e-2) on these scenes; atomics make the sum order -- here and in the reference -- vary from run to run.
    dev_t = np.abs(o32["poses"][:, :3] - o["poses"][:, :3]).max()
    dev_r = quat_angle(o32["poses"][:, 3:], o["poses"][:, 3:]).max()
    dev_d = np.abs(o32["patches"][:, 2, 0, 0] - o["patches"][:, 2, 0, 0]).max()
    return o, max(2e-4, 3 * dev_t), max(2e-4, 3 * dev_r), max(2e-3 * max(1.0, np.abs(o["patches"][:, 2, 0, 0]).max()), 3 * dev_d)


def _check_global(poses, patches, o, tol_t, tol_r, tol_d):
    assert np.abs(poses[:, :3] - o["poses"][:, :3]).max() < tol_t
    assert quat_angle(poses[:, 3:], o["poses"][:, 3:]).max() < tol_r
    assert np.abs(patches[:, 2, 0, 0] - o["patches"][:, 2, 0, 0]).max() < tol_d
    assert np.array_equal(patches[:, 2], np.broadcast_to(patches[:, 2, :1, :1], patches[:, 2].shape))


@pytest.mark.parametrize("kw,t0", [(dict(seed=11, n=40, M=8, edges="tracker"), 1), (dict(seed=12, n=24, M=6, edges="all"), 1),
                                   (dict(seed=13, n=64, M=8, edges="tracker"), 20), (dict(seed=15, n=48, M=6, edges="global"), 1)])
def test_global_ba_vs_oracle(syn, kw, t0):
    """more than 16 free poses: cuda_ba.forward takes the dense-in-HBM path (wv_ba_global_*)"""
    s = syn.make_ba_scene(**kw)
    t1 = int(s["t1"])
    assert t1 - t0 > 16
    o, *tols = _global_tol(s, t0, t1, 2)
    poses, patches = run_ba(s, t0, t1, 2)
    _check_global(poses, patches, o, *tols)
    assert np.array_equal(poses[:t0], s["poses"][:t0])  # fixed poses untouched


def test_global_ba_depth_prior_and_general_graph(syn):
    s = syn.make_ba_scene(seed=14, n=30, M=5, edges="all")
    rng = np.random.default_rng(1)
    dup = rng.integers(0, len(s["kk"]), 60)
    for k in ("ii", "jj", "kk", "target", "weight"):
        s[k] = np.concatenate([s[k], s[k][dup]])
    s["ii"][-10:] = (s["ii"][-10:] + 1) % 30  # some patches seen from two source frames (per-edge atomics path)
    perm = rng.permutation(len(s["kk"]))
    for k in ("ii", "jj", "kk", "target", "weight"):
        s[k] = np.ascontiguousarray(s[k][perm])
    o, *tols = _global_tol(s, 1, 30, 2, patches_est=s["patches_est"], mu=2.0)
    poses, patches = run_ba(s, 1, 30, 2, patches_est=s["patches_est"], mu=2.0)
    _check_global(poses, patches, o, *tols)


def test_global_pieces_match_windowed_kernel_and_shard_sum(syn):
    """the same 10-pose window through the global pieces == the windowed kernel (rounding), and the partial systems / depth
    updates of 3 patch shards add up to the unsharded ones"""
    _, cb = pkg().native_modules()
    s = syn.make_ba_scene(seed=2, n=40, M=32, edges="tracker")
    t1 = int(s["t1"]); t0 = t1 - 10
    D = {k: dev(s[k]) for k in ("intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")}
    args = lambda P, Q: (P, Q, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1)
    wp, wq = run_ba(s, t0, t1, 1)
    poses, patches = dev(s["poses"].copy()), dev(s["patches"].copy())
    Sy = cb.global_build(*args(poses, patches), 0, 1, True)
    parts = [cb.global_build(*args(poses, patches), r, 3, False) for r in range(3)]
    ssum = parts[0] + parts[1] + parts[2]
    scale = Sy.abs().max().item()
    assert (ssum - Sy).abs().max().item() <= 1e-5 * scale
    n6 = 60
    S = Sy[: n6 * n6].view(n6, n6)
    assert torch.count_nonzero(torch.triu(S, 1)).item() == 0  # only the lower triangle is written
    dX = cb.global_solve(Sy, t1 - t0)
    dZ = cb.global_dz(*args(poses, patches), 0, 1, dX)
    dzp = [cb.global_dz(*args(poses, patches), r, 3, dX) for r in range(3)]
    assert torch.equal(dzp[0] + dzp[1] + dzp[2], dZ)  # disjoint supports: exact
    assert all(torch.count_nonzero(z).item() > 0 for z in dzp)
    cb.global_retract(poses, patches, D["ii"].numel(), t0, t1, dX, dZ)
    torch.cuda.synchronize()
    gp, gq = poses.cpu().numpy(), patches.cpu().numpy()
    assert np.abs(gp[:, :3] - wp[:, :3]).max() < 1e-4 and quat_angle(gp[:, 3:], wp[:, 3:]).max() < 1e-4
    assert np.abs(gq[:, 2] - wq[:, 2]).max() < 1e-3 * max(1.0, np.abs(wq[:, 2]).max())


def test_global_ba_sharded_wrapper_single_rank(syn):
    """fastba.BA_sharded without a process group == cuda_ba.forward on the global path"""
    fb = importlib.import_module(PKG + ".fastba")
    s = syn.make_ba_scene(seed=11, n=40, M=8, edges="tracker")
    t1 = int(s["t1"])
    o, *tols = _global_tol(s, 1, t1, 2)
    poses, patches = dev(s["poses"].copy()), dev(s["patches"].copy())
    fb.BA_sharded(poses, patches, dev(s["intrinsics"]), dev(s["target"]), dev(s["weight"]), dev(s["lmbda"]), dev(s["ii"]),
                  dev(s["jj"]), dev(s["kk"]), 1, t1, 2)
    torch.cuda.synchronize()
    _check_global(poses.cpu().numpy(), patches.cpu().numpy(), o, *tols)


def test_global_ba_not_spd_raises(syn):
    _, cb = pkg().native_modules()
    s = syn.make_ba_scene(seed=11, n=40, M=8, edges="tracker")
    s["weight"] = -np.abs(s["weight"]) * 50.0  # negative weights: S is not positive-definite
    with pytest.raises(RuntimeError, match="positive-definite"):
        run_ba(s, 1, int(s["t1"]), 2)
    s2 = syn.make_ba_scene(seed=11, n=40, M=8, edges="tracker")
    run_ba(s2, 1, int(s2["t1"]), 1)  # the extension still works afterwards


# ------------------------------------------------------------------------------------------------- corr at full size
@pytest.mark.parametrize("M", [96, 384])
def test_corr_full_size_properties(syn, M):
    """BASELINE configs[1] size (M = 96, E = 45 312) and default.yaml's M = 384 (E = 181 248), both pyramid levels, fp16 maps:
    tiled == generic kernel to one fp16 rounding; a 256-edge sample == fp64 oracle; and size-independent exact properties:
    scaling the map by 2 doubles the output bit for bit (where the fp16 result is a normal number), permuting the edges permutes the output bit for bit, edges whose
    window lies wholly outside the map give exact zeros."""
    cc, _ = pkg().native_modules()
    w = syn.make_corr_workload(seed=21, M=M)
    rng = np.random.default_rng(5)
    E = w["coords"].shape[1]
    g = dev(w["gmap"], torch.float16)
    ii, jj = dev(w["ii1"]), dev(w["jj1"])
    perm = torch.from_numpy(rng.permutation(E)).cuda()
    for lvl, (fm, div) in enumerate(((w["fmap0"], 1.0), (w["fmap1"], 4.0))):
        f = dev(fm, torch.float16)
        co = dev((w["coords"] / div).astype(np.float32))
        H, W = fm.shape[3], fm.shape[4]
        with _tiled(True):
            out, = cc.forward(g, f, co, ii, jj, 3)
            out2, = cc.forward(g, f * 2, co, ii, jj, 3)
            outp, = cc.forward(g, f, co[:, perm].contiguous(), ii[perm].contiguous(), jj[perm].contiguous(), 3)
        with _tiled(False):
            gen, = cc.forward(g, f, co, ii, jj, 3)
        torch.cuda.synchronize()
        scale = max(1.0, out.float().abs().max().item())
        assert (out.float() - gen.float()).abs().max().item() <= 2.0 ** -9 * scale, f"level {lvl}"
        normal = out.abs() >= 2.0 ** -13  # fp16 results in the subnormal range round on a fixed grid: scaling is exact only above it
        assert torch.equal(out2[normal], (out * 2)[normal]) and normal.float().mean().item() > 0.8, f"level {lvl}: linearity"
        assert (out2.float() - 2 * out.float()).abs().max().item() <= 2.0 ** -23, f"level {lvl}: linearity (subnormal results)"
        assert torch.equal(outp, out[:, perm]), f"level {lvl}: permutation"
        c = co[0]  # [E,2,3,3]
        outside = ((c[:, 0] < -5).all(-1).all(-1) | (c[:, 0] > W + 4).all(-1).all(-1) | (c[:, 1] < -5).all(-1).all(-1) | (c[:, 1] > H + 4).all(-1).all(-1))
        assert outside.any() and out[0, outside].abs().max().item() == 0.0, f"level {lvl}: out-of-map edges"
        pick = np.sort(rng.choice(E, 256, replace=False))
        g16 = torch.from_numpy(w["gmap"]).half().float().numpy()
        f16 = torch.from_numpy(fm).half().float().numpy()
        truth = oracle.corr_forward(g16, f16, (w["coords"][:, pick] / div).astype(np.float32), w["ii1"][pick], w["jj1"][pick], 3, mode=oracle.F64)
        got = out[:, torch.from_numpy(pick).cuda()].float().cpu().numpy()
        assert np.abs(got - truth).max() <= 2.0 ** -10 * max(1.0, np.abs(truth).max()), f"level {lvl}: oracle sample"


def test_planned_ba_detects_a_plan_built_for_other_edges(syn):
    """C ABI: wv_ba_forward_planned on a workspace whose plan belongs to a different kk must say so (WV_STATUS_STALE_PLAN),
    and kk values beyond the patch buffer on a reused plan must be flagged, not dereferenced"""
    lib = pkg().capi()
    s = syn.make_ba_scene(seed=2, n=20, M=16, edges="tracker")
    t1 = int(s["t1"]); t0 = t1 - 10
    D = {k: dev(s[k]) for k in ("poses", "patches", "intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")}
    E, groups = D["kk"].numel(), 1 << 16
    n_poses, n_patches = D["poses"].shape[0], D["patches"].shape[0]
    lib.wv_ba_workspace_bytes.restype = ctypes.c_size_t
    lib.wv_ba_workspace_bytes.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.c_int]
    wsb = lib.wv_ba_workspace_bytes(E, groups, t1 - t0)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    ix, jx = torch.empty_like(D["kk"]), torch.empty_like(D["kk"])
    P = ctypes.c_void_p

    def planned(kk_plan, kk_ba, n_patches_arg):
        status.zero_()
        rc = lib.wv_neighbors(P(kk_plan.data_ptr()), P(D["jj"].data_ptr()), ctypes.c_int64(E), ctypes.c_int64(groups), P(ix.data_ptr()),
                              P(jx.data_ptr()), P(ws.data_ptr()), ctypes.c_size_t(wsb), P(status.data_ptr()), None)
        assert rc == 0
        poses, patches = D["poses"].clone(), D["patches"].clone()
        rc = lib.wv_ba_forward_planned(P(poses.data_ptr()), P(patches.data_ptr()), P(D["intrinsics"].data_ptr()), P(D["target"].data_ptr()),
                                       P(D["weight"].data_ptr()), P(D["lmbda"].data_ptr()), P(D["ii"].data_ptr()), P(D["jj"].data_ptr()),
                                       P(kk_ba.data_ptr()), ctypes.c_int64(E), ctypes.c_int64(n_poses), ctypes.c_int64(n_patches_arg),
                                       ctypes.c_int(3), ctypes.c_int(t0), ctypes.c_int(t1), ctypes.c_int(2), None, ctypes.c_float(2.0),
                                       P(ws.data_ptr()), ctypes.c_size_t(wsb), ctypes.c_int64(groups), P(status.data_ptr()), None)
        assert rc == 0
        torch.cuda.synchronize()
        return int(status.item()), poses, patches

    st, poses, _ = planned(D["kk"], D["kk"], n_patches)
    assert st == 0 and not torch.equal(poses, D["poses"])
    other = D["kk"].clone(); other[::7] = (other[::7] + 3) % n_patches
    st, poses, patches = planned(D["kk"], other, n_patches)
    assert st & 8                      # WV_STATUS_STALE_PLAN ...
    assert torch.equal(poses, D["poses"]) and torch.equal(patches, D["patches"])   # ... and nothing was applied
    st, _, patches = planned(D["kk"], D["kk"], n_patches // 2)   # half the patch buffer: ids beyond it are flagged ...
    assert st & 1                      # WV_STATUS_BAD_INDEX
    assert torch.equal(patches[n_patches // 2:], D["patches"][n_patches // 2:])  # ... and never touched


def test_forward_repeats_with_its_own_plan_when_the_cached_plan_is_stale(syn):
    """cuda_ba.neighbors(kk, jj) leaves an edge plan behind for the BA that follows on the same tensors (one update() of the tracker).
    If the tensor at that address holds OTHER ids by the time forward() runs -- same pointer, same _version: what a caching allocator
    can produce after the first tensor died -- the plan is stale: the call must come back with the right answer, not a corrupted state
    and an exception (the kernel verifies every edge against the plan, stops before applying, the binding repeats)."""
    _, cb = pkg().native_modules()
    s = syn.make_ba_scene(seed=2, n=20, M=16, edges="tracker")
    t1 = int(s["t1"]); t0 = t1 - 10
    D = {k: dev(s[k]) for k in ("intrinsics", "target", "weight", "lmbda", "ii", "jj")}
    kk_true = dev(s["kk"])
    n_patches = s["patches"].shape[0]
    wrong = kk_true.clone(); wrong[::7] = (wrong[::7] + 3) % n_patches
    kk = wrong.clone()
    cb.neighbors(kk, D["jj"])                                  # the plan left behind is for the WRONG ids ...
    v = kk._version
    kk.data.copy_(kk_true)                                     # ... same storage, same version counter, other contents (`.data` writes
    assert kk._version == v and torch.equal(kk, kk_true)       # do not bump the version): the cache key still matches
    poses, patches = dev(s["poses"].copy()), dev(s["patches"].copy())
    cb.forward(poses, patches, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], kk, t0, t1, 2)
    wp, wq = run_ba(s, t0, t1, 2)                              # the same call without any cached plan
    assert np.array_equal(poses.cpu().numpy(), wp) and np.array_equal(patches.cpu().numpy(), wq)


def test_corr_index_error_surfaces_through_the_shared_status_word(syn):
    """a bad index in a correlation call zeroes that edge's rows and flags the library's status word -- the SAME word cuda_ba reads
    (each extension module is its own shared object; round 1 kept one word per module, so nobody ever saw this flag)"""
    cc, cb = pkg().native_modules()
    w = syn.make_corr_workload(seed=1, M=8, pmem=12, n_frames=14, ht=24, wd=32, C=128, lifetime=6, removal=9)
    g, f0 = dev(w["gmap"], torch.float16), dev(w["fmap0"], torch.float16)
    co, ii1, jj1 = dev(w["coords"]), dev(w["ii1"]), dev(w["jj1"])
    cb.check_status(co)                                         # start clean
    bad = jj1.clone(); bad[3] = 10 ** 6
    out, = cc.forward(g, f0, co, ii1, bad, 3)
    assert out[0, 3].abs().max().item() == 0.0
    with pytest.raises(IndexError):
        cb.check_status(co)
    cb.check_status(co)                                         # reading cleared it


# ------------------------------------------------------------------------------------------------- fused projective ops (N1 / N4)
def _pops_dev(g, name):
    return [dev(g[f"{name}_{k}"]) for k in ("poses", "patches", "intrinsics", "ii", "jj", "kk")]


@pytest.mark.parametrize("name", ["trk", "glb"])
def test_pops_matches_reference_python(golden, name):
    """projective_ops mirror (one kernel per call) vs outputs of the reference's own dpvo/projective_ops.py + lietorch CPU ops.
    tolerance (stated): 2e-4 px on coordinates of magnitude ~150 px, 2e-6 relative on 3-D points; validity flags exact"""
    pops = importlib.import_module(PKG + ".projective_ops")
    g = golden["pops_ref"]
    a = _pops_dev(g, name)
    a[0], a[1], a[2] = a[0][None], a[1][None], a[2][None]  # the reference's [1, N, ...] layout
    co = pops.transform(*a)
    assert tuple(co.shape) == (1,) + g[f"{name}_transform"].shape
    assert np.abs(co[0].cpu().numpy() - g[f"{name}_transform"]).max() < 2e-4
    cd, val = pops.transform(*a, depth=True, valid=True)
    assert np.abs(cd[0].cpu().numpy() - g[f"{name}_transform_depth"]).max() < 2e-4
    assert np.array_equal(val[0].cpu().numpy(), g[f"{name}_valid"])
    assert np.abs(pops.transform(*a, tonly=True)[0].cpu().numpy() - g[f"{name}_transform_tonly"]).max() < 2e-4
    for beta in (0.3, 0.5):
        assert np.abs(pops.flow_mag(*a, beta=beta)[0].cpu().numpy() - g[f"{name}_flow_mag_{beta}"]).max() < 2e-4
    pc = pops.point_cloud(a[0], a[1], a[2], dev(g[f"{name}_ix"]))[0].cpu().numpy()
    assert np.abs(pc - g[f"{name}_point_cloud"]).max() < 2e-6 * max(1.0, np.abs(g[f"{name}_point_cloud"]).max())
    # jacobian=True (projective_ops.py:71-99): centre-pixel Jacobians; entries reach ~1e3 (fx X d^2 with small Z), so the stated
    # tolerance is relative to each tensor's peak: 2e-5
    xj, vj, (Ji, Jj, Jz) = pops.transform(*a, jacobian=True)
    assert np.abs(xj[0].cpu().numpy() - g[f"{name}_jac_x1"]).max() < 2e-4
    assert np.array_equal(vj[0].cpu().numpy(), g[f"{name}_jac_valid"])
    for got, key in ((Ji, "Ji"), (Jj, "Jj"), (Jz, "Jz")):
        want = g[f"{name}_jac_{key}"]
        assert tuple(got.shape) == (1,) + want.shape
        assert np.abs(got[0].cpu().numpy() - want).max() <= 2e-5 * max(1.0, np.abs(want).max()), key

    class FakeSE3:  # anything with .data, like lietorch's SE3
        data = a[0]
    assert torch.equal(pops.transform(FakeSE3(), *a[1:]), co)


@pytest.mark.parametrize("name", ["trk", "glb"])
def test_frame_distance_and_edges_match_reference(golden, name):
    pops = importlib.import_module(PKG + ".projective_ops")
    g = golden["pops_ref"]
    a = _pops_dev(g, name)
    Dref = g[f"{name}_frame_distance"]
    n = Dref.shape[0]
    M = 6 if name == "trk" else 4
    D = pops.frame_distance(a[0], a[1], a[2], n, M, beta=0.5).cpu().numpy()
    assert np.abs(D - Dref).max() < 2e-4 * max(1.0, Dref.max()) and np.array_equal(D, D.T) and np.all(np.diag(D) == 0)
    ei, ej = pops.distance_based_edges(a[0], a[1], a[2], n, M, float(g[f"{name}_thresh"]), beta=0.5)
    assert ei == g[f"{name}_edges_ii"].tolist() and ej == g[f"{name}_edges_jj"].tolist()   # bit-exact edge list, reference order


def test_pops_full_size_vs_oracle_and_bad_index(syn):
    """tracker-size edge list (E = 45 312): transform / flow_mag vs the fp64 oracle; out-of-range index raises"""
    pops = importlib.import_module(PKG + ".projective_ops")
    s = syn.make_ba_scene(seed=2, n=40, M=96, edges="tracker")
    a = [dev(s[k]) for k in ("poses", "patches", "intrinsics", "ii", "jj", "kk")]
    co = pops.transform(*a)[0].cpu().numpy()
    ref = oracle.pops_transform(*[s[k] for k in ("poses", "patches", "intrinsics", "ii", "jj", "kk")])[0]
    far = np.abs(ref) > 1e4  # points projected through d = 1 / 0.1 from behind the camera: compare relatively
    assert np.abs(co - ref)[~far].max() < 1e-3 and np.all(np.abs(co - ref)[far] <= 1e-5 * np.abs(ref)[far])
    fm = pops.flow_mag(*a, beta=0.5)[0].cpu().numpy()
    fref = oracle.pops_flow_mag(*[s[k] for k in ("poses", "patches", "intrinsics", "ii", "jj", "kk")], beta=0.5)[0]
    assert np.abs(fm - fref).max() < 1e-3 * max(1.0, np.abs(fref).max() * 1e-2)
    bad = a[4].clone(); bad[3] = 10 ** 6
    os.environ["WVB200_CHECK_INDICES"] = "1"
    try:
        with pytest.raises(IndexError):
            pops.transform(a[0], a[1], a[2], a[3], bad, a[5])
    finally:
        del os.environ["WVB200_CHECK_INDICES"]


def test_global_ba_dense_patches_and_heavy_patches(syn):
    """every patch observed by every frame.  60 observations per patch go through the per-warp entry list in shared memory (2 lane
    chunks per patch); 399 observations per patch (more than the 384-entry list holds: round 1 refused that) go through the heavy
    pass with the entry lists in global memory -- the reference has no limit (a camera turning on the spot under distance-based
    edges sees a patch from every frame).  Both against the fp64 oracle; a mixed scene (heavy and light patches side by side) too."""
    s = syn.make_ba_scene(seed=16, n=60, M=2, edges="all", pose_noise=0.005)
    o, *tols = _global_tol(s, 1, 60, 1)
    poses, patches = run_ba(s, 1, 60, 1)
    _check_global(poses, patches, o, *tols)
    big = syn.make_ba_scene(seed=17, n=400, M=1, edges="all", pose_noise=0.002)
    o, *tols = _global_tol(big, 1, 400, 1)
    poses, patches = run_ba(big, 1, 400, 1)
    _check_global(poses, patches, o, *tols)
    assert np.abs(patches[:, 2, 0, 0] - big["patches"][:, 2, 0, 0]).max() > 1e-4     # the depths did move
    # mixed: the global pattern plus three patches that every frame observes (the edge list stays in the generator's order: the
    # heavy patches' edges come last, not grouped with their neighbours')
    n = 420
    mix = syn.make_ba_scene(seed=18, n=n, M=2, edges="global", pose_noise=0.003, heavy_patches=(7, 401, 838))
    assert np.bincount(mix["kk"]).max() == n - 1
    o, *tols = _global_tol(mix, 1, n, 1)
    poses, patches = run_ba(mix, 1, n, 1)
    _check_global(poses, patches, o, *tols)


# ------------------------------------------------------------------------------------------------- SoftAgg (group plan + fused scatter softmax / sum)
def _keys(syn, kind, **kw):
    s = syn.make_ba_scene(**kw)
    return (s["kk"] if kind == "kk" else s["ii"] * 12345 + s["jj"]).astype(np.int64)   # dpvo/net.py:88-89


@pytest.fixture(params=["coop", "bitmap"])
def plan_path(request):
    """both group-plan implementations: the cooperative counting sort shared with the BA (default) and the bitmap + popcount-rank
    path that serves key ranges beyond 2^22"""
    if request.param == "bitmap":
        os.environ["WVB200_AGG_BITMAP_PLAN"] = "1"
    yield request.param
    os.environ.pop("WVB200_AGG_BITMAP_PLAN", None)


@pytest.mark.parametrize("kind", ["kk", "ij"])
def test_group_plan_bit_exact_vs_torch_unique(syn, kind, plan_path):
    agg = pkg().native_agg()
    for kw in (dict(seed=5, n=12, M=6, edges="tracker", lifetime=4, removal=6), dict(seed=2, n=40, M=96, edges="tracker")):
        ix = dev(_keys(syn, kind, **kw))
        tu, tj = torch.unique(ix, return_inverse=True)          # the reference's call, dpvo/blocks.py:41
        for hint in ((0, -1), (int(ix.min()), int(ix.max())), (int(ix.min()) - 3, int(ix.max()) + 1000)):
            jx, order, offsets, ng, G = agg.group_plan(ix, *hint)
            assert G == tu.numel() == int(ng[0].item()) and torch.equal(jx, tj)
            order, offsets = order.cpu().numpy(), offsets.cpu().numpy()
            j = tj.cpu().numpy()
            assert offsets[0] == 0 and offsets[G] == len(j) and np.array_equal(np.diff(offsets[:G + 1]), np.bincount(j, minlength=G))
            assert np.array_equal(order, np.argsort(j, kind="stable"))   # grouped, ascending edge index inside a group


def test_group_plan_edge_cases(syn, plan_path):
    agg = pkg().native_agg()
    jx, order, offsets, ng, G = agg.group_plan(torch.zeros(0, dtype=torch.long, device="cuda"))
    assert G == 0 and jx.numel() == 0
    ix = torch.tensor([5, -7, 5, 900000, -7, 5], device="cuda")
    jx, order, offsets, ng, G = agg.group_plan(ix)
    assert G == 3 and jx.tolist() == [1, 0, 1, 2, 0, 1] and order.tolist() == [1, 4, 0, 2, 5, 3] and offsets.tolist()[:4] == [0, 2, 5, 6]
    big = torch.arange(0, 5000, device="cuda").repeat_interleave(40)[torch.randperm(200000, device="cuda")]   # groups of 40 > one warp
    jx, order, offsets, ng, G = agg.group_plan(big)
    assert G == 5000 and torch.equal(jx, big) and np.array_equal(order.cpu().numpy(), np.argsort(big.cpu().numpy(), kind="stable"))
    with pytest.raises(IndexError):
        agg.group_plan(ix, 0, 100)                              # keys outside the promised bounds
    assert agg.group_plan(ix)[4] == 3                           # the failure left nothing behind
    with pytest.raises(RuntimeError):
        agg.group_plan(torch.tensor([0, 1 << 40], device="cuda"))
    wide = torch.tensor([3, (1 << 27) + 5, 3, 17, (1 << 27) + 5], device="cuda")     # range > 2^22: always the bitmap path
    jx, order, offsets, ng, G = agg.group_plan(wide)
    assert G == 3 and jx.tolist() == [0, 2, 0, 1, 2] and order.tolist() == [0, 2, 3, 1, 4]


@pytest.mark.parametrize("dtype,D,tol", [(torch.float16, 384, 2.0 ** -10), (torch.float32, 384, 2e-6), (torch.float16, 20, 2.0 ** -10),
                                         (torch.float32, 7, 2e-6)])
def test_soft_agg_vs_oracle(syn, dtype, D, tol):
    """tolerance: fp32 arithmetic + ONE rounding to the output dtype -> half an ulp of the result scale (2^-11 relative for fp16;
    the reference's own fp16 evaluation is ~10x further from the exact value, see test_soft_agg_oracle_properties)"""
    blocks = importlib.import_module(PKG + ".blocks")
    rng = np.random.default_rng(31)
    ix = _keys(syn, "ij", seed=5, n=12, M=6, edges="tracker", lifetime=4, removal=6)
    E = len(ix)
    B = 2
    f = dev(rng.standard_normal((B, E, D)).astype(np.float32), dtype)
    g = dev((3 * rng.standard_normal((B, E, D))).astype(np.float32), dtype)
    jx, plan = blocks.group_plan(dev(ix))
    y = blocks.scatter_softmax_sum(f, g, plan)
    y2 = blocks.scatter_softmax_sum(f, g, plan)
    truth, tj = oracle.soft_agg(f.float().cpu().numpy().astype(np.float64), g.float().cpu().numpy().astype(np.float64), ix)
    assert y.dtype == dtype and tuple(y.shape) == truth.shape and np.array_equal(jx.cpu().numpy(), tj)
    assert torch.equal(y, y2)                                   # fixed summation order
    scale = max(1.0, np.abs(truth).max())
    assert np.abs(y.float().cpu().numpy() - truth).max() <= tol * scale


def test_soft_agg_module_matches_composed_reference_form(syn):
    """blocks.SoftAgg (expand=True / False, autocast like the tracker) against the same module evaluated with torch ops in the
    reference's order: unique -> softmax per group -> weighted sum -> h -> expand (dpvo/blocks.py:40-48)"""
    blocks = importlib.import_module(PKG + ".blocks")
    torch.manual_seed(0)
    ix = dev(_keys(syn, "kk", seed=6, n=12, M=6, edges="tracker", lifetime=4, removal=6))
    E, D = ix.numel(), 384
    x = torch.randn(1, E, D, device="cuda")
    for expand in (True, False):
        m = blocks.SoftAgg(D, expand=expand).cuda()
        with torch.no_grad(), torch.autocast("cuda", enabled=True):
            out = m(x, ix)
            _, jx = torch.unique(ix, return_inverse=True)
            y = blocks._soft_agg_autograd(m.f(x).float(), m.g(x).float(), jx, int(jx.max()) + 1)
            ref = m.h(y.half())[:, jx] if expand else m.h(y.half())
        assert out.shape == ref.shape and out.dtype == torch.float16
        assert (out.float() - ref.float()).abs().max() <= 4e-3 * max(1.0, ref.float().abs().max().item())
    # gradients flow through the composed form (training path)
    m = blocks.SoftAgg(D).cuda()
    xg = x.clone().requires_grad_(True)
    m(xg, ix).sum().backward()
    assert xg.grad is not None and torch.isfinite(xg.grad).all() and m.g.weight.grad.abs().sum() > 0


def test_soft_agg_full_size_properties(syn, plan_path):
    """E = 45 312, D = 384 fp16 (the tracker's update operator): size-independent properties + an oracle sample"""
    blocks = importlib.import_module(PKG + ".blocks")
    rng = np.random.default_rng(33)
    for kind in ("kk", "ij"):
        ixn = _keys(syn, kind, seed=2, n=40, M=96, edges="tracker")
        ix = dev(ixn)
        E, D = len(ixn), 384
        f = torch.randn(1, E, D, device="cuda").half(); g = (3 * torch.randn(1, E, D, device="cuda")).half()
        jx, plan = blocks.group_plan(ix)
        y = blocks.scatter_softmax_sum(f, g, plan)
        G = plan[3]
        assert tuple(y.shape) == (1, G, D) and torch.isfinite(y).all()
        # weights of a group sum to one: constant f -> constant y (exactly representable constant)
        yc = blocks.scatter_softmax_sum(torch.full_like(f, 0.75), g, plan)
        assert (yc.float() - 0.75).abs().max() <= 2.0 ** -11
        # shift invariance in g (exact shift by a representable constant per channel is not exact in fp16; use linearity in f instead)
        y2 = blocks.scatter_softmax_sum(2 * f, g, plan)          # linear in f: exact in fp32, identical roundings above fp16 subnormals
        assert (y2.float() - 2 * y.float()).abs().max() <= 2.0 ** -23
        # a bound on G instead of the host read of G: same rows, zeros beyond, nothing waits for the device
        lo_k, hi_k = int(ixn.min()), int(ixn.max())
        jb, planb = blocks.group_plan(ix, lo_k, hi_k, g_capacity=G + 37)
        yb = blocks.scatter_softmax_sum(f, g, planb)
        assert torch.equal(jb, jx) and tuple(yb.shape) == (1, G + 37, D) and torch.equal(yb[:, :G], y) and (yb[:, G:] == 0).all()
        # convexity: every output lies inside the range of its group's f values
        idx = jx.view(1, -1, 1).expand(1, E, D)
        lo = torch.full((1, G, D), float("inf"), device="cuda").scatter_reduce(1, idx, f.float(), "amin")
        hi = torch.full((1, G, D), float("-inf"), device="cuda").scatter_reduce(1, idx, f.float(), "amax")
        assert (y.float() >= lo - 1e-3).all() and (y.float() <= hi + 1e-3).all()
        # permuting the edges permutes nothing in y (groups are sets) up to the fp32 summation order
        perm = torch.randperm(E, device="cuda")
        jp, planp = blocks.group_plan(ix[perm])
        yp = blocks.scatter_softmax_sum(f[:, perm], g[:, perm], planp)
        assert (yp.float() - y.float()).abs().max() <= 2.0 ** -9 * max(1.0, y.float().abs().max().item())
        # oracle on a sample of groups
        sel = rng.choice(G, 24, replace=False)
        jn = jx.cpu().numpy()
        keep = np.isin(jn, sel)
        truth, _ = oracle.soft_agg(f[:, keep].float().cpu().numpy().astype(np.float64), g[:, keep].float().cpu().numpy().astype(np.float64), ixn[keep])
        got = y[:, np.sort(sel)].float().cpu().numpy()
        assert np.abs(got - truth).max() <= 2.0 ** -10 * max(1.0, np.abs(truth).max())


# ------------------------------------------------------------------------------------------------- the reference's own CUDA build
def test_full_size_parity_vs_the_reference_cuda_build(syn):
    """oracle/_ref/refgpu_*.so = the reference's own correlation_kernel.cu / ba_cuda.cu / ba.cpp compiled UNEDITED for sm_100
    (oracle/build_ref_gpu.py; ~10 min of nvcc, so it is built on demand and this test skips when it is absent).  Same inputs as
    bench.py's step, full size: corr within the reference's own fp16-accumulation error, neighbors identical, BA within the fp32
    reassociation tolerance of its atomics."""
    import build_ref
    rc, rb = build_ref.load("refgpu_cuda_corr"), build_ref.load("refgpu_cuda_ba")
    if rc is None or rb is None:
        pytest.skip("oracle/_ref/refgpu_*.so not built")
    cc, cb = pkg().native_modules()
    altcorr = importlib.import_module(PKG + ".altcorr")
    w = syn.make_corr_workload(seed=1)
    s = syn.make_ba_scene(seed=2, n=40, M=96, edges="tracker")
    E = w["E"]
    g, f0, f1 = dev(w["gmap"], torch.float16), dev(w["fmap0"], torch.float16), dev(w["fmap1"], torch.float16)
    co, ii1, jj1 = dev(w["coords"]), dev(w["ii1"]), dev(w["jj1"])
    ref = torch.stack([rc.forward(g, f0, co / 1, ii1, jj1, 3)[0], rc.forward(g, f1, co / 4, ii1, jj1, 3)[0]], -1).view(1, E, -1)
    ours = altcorr.corr_stacked(g, (f0, f1), co, ii1, jj1, 3)
    scale = max(1.0, ref.float().abs().max().item())
    # the reference sums 128 fp16 products in fp16 (correlation_kernel.cu:117-133): ~sqrt(128) * 2^-11 relative, measured 4e-3
    assert (ref.float() - ours.float()).abs().max().item() <= 1e-2 * scale
    ix, jx = cb.neighbors(dev(s["kk"]), dev(s["jj"]))
    rix, rjx = rb.neighbors(dev(s["kk"]), dev(s["jj"]))
    assert torch.equal(ix, rix.to(ix.device)) and torch.equal(jx, rjx.to(jx.device))
    t0, t1 = int(s["t1"]) - 10, int(s["t1"])
    outs = []
    for mod in (rb, cb):
        po, pa = dev(s["poses"].copy()), dev(s["patches"].copy())
        mod.forward(po, pa, dev(s["intrinsics"]), dev(s["target"]), dev(s["weight"]), dev(s["lmbda"]), dev(s["ii"]), dev(s["jj"]), dev(s["kk"]), t0, t1, 2)
        outs.append((po, pa))
    assert (outs[0][0] - outs[1][0]).abs().max().item() < 1e-4 and (outs[0][1][:, 2] - outs[1][1][:, 2]).abs().max().item() < 1e-3


def _cam_centres(poses):
    """world positions of the cameras from world->camera poses [t, q(xyzw)], fp64"""
    p = np.asarray(poses, np.float64)
    t, q = p[:, :3], p[:, 3:] / np.linalg.norm(p[:, 3:], axis=1, keepdims=True)
    x, y, z, w = q.T
    R = np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                  np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                  np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], -2)
    return -np.einsum("nji,nj->ni", R, t)


def _sim3_max_diff(a, b):
    """max distance between two sets of camera centres after the least-squares Sim(3) alignment of a onto b (Umeyama): removes the
    monocular scale gauge, which only the +1 damping holds (ba_cuda.cu:517-518)"""
    mu_a, mu_b = a.mean(0), b.mean(0)
    xa, xb = a - mu_a, b - mu_b
    U, D, Vt = np.linalg.svd(xb.T @ xa / len(a))
    S = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        S[2, 2] = -1
    R = U @ S @ Vt
    sc = np.trace(np.diag(D) @ S) / ((xa ** 2).sum() / len(a))
    return float(np.linalg.norm(sc * xa @ R.T + mu_b - b, axis=1).max()), float(sc)


@pytest.mark.parametrize("solver", ["band", "dense"])
def test_global_ba_at_512_frames_vs_the_reference_cuda_build_gauge_free(syn, solver, monkeypatch):
    """global BA at scale against the reference's own CUDA build (refgpu_cuda_ba: ba_cuda.cu unedited for
    sm_100), 511 free poses, the edge pattern of the reference's global_bundle_adjustment (dpvo/dpvo.py:448-471).  Compared gauge-free:
    (a) the weighted reprojection cost after 2 iterations, (b) the camera centres after Sim(3) alignment -- beside the reference's OWN
    run-to-run spread (its atomics reorder the sums), which is the yardstick: this build must be no further from a reference run than
    a few times what two reference runs are from each other."""
    import build_ref
    rb = build_ref.load("refgpu_cuda_ba")
    if rb is None:
        pytest.skip("oracle/_ref/refgpu_cuda_ba.so not built")
    if solver == "dense":
        monkeypatch.setenv("WVB200_GLOBAL_DENSE", "1")
    _, cb = pkg().native_modules()
    s = syn.make_ba_scene(seed=31, n=512, M=16, edges="global", pose_noise=0.005)
    t0, t1 = 1, 512
    D = {k: dev(s[k]) for k in ("intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")}

    def run(mod):
        po, pa = dev(s["poses"].copy()), dev(s["patches"].copy())
        mod.forward(po, pa, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, 2)
        torch.cuda.synchronize()
        co = cb.reproject(po, pa, D["intrinsics"], D["ii"], D["jj"], D["kk"])[0, :, :, 1, 1].double()
        r = D["target"].double() - co
        ok = torch.isfinite(r).all(-1) & (r.norm(dim=-1) < 128)
        return _cam_centres(po.cpu().numpy()), pa[:, 2, 0, 0].double().cpu().numpy(), float((D["weight"].double() * r * r)[ok].sum())

    c_r1, d_r1, cost_r1 = run(rb)
    c_r2, d_r2, cost_r2 = run(rb)
    c_o, d_o, cost_o = run(cb)
    c0, _, cost0 = _cam_centres(s["poses"]), None, None
    spread, _ = _sim3_max_diff(c_r2, c_r1)
    diff, sc = _sim3_max_diff(c_o, c_r1)
    moved, _ = _sim3_max_diff(c0, c_r1)
    extent = float(np.linalg.norm(c_r1.max(0) - c_r1.min(0)))
    msg = dict(solver=solver, cost_ref=cost_r1, cost_ref_again=cost_r2, cost_ours=cost_o, spread_ref_vs_ref=spread, ours_vs_ref=diff, scale=sc,
               input_vs_ref=moved, extent=extent)
    print(msg)
    assert abs(cost_o - cost_r1) <= max(2e-3 * cost_r1, 5 * abs(cost_r2 - cost_r1)), msg
    assert diff <= max(5 * spread, 2e-4 * extent), msg           # and far below how much the BA moved the cameras:
    assert diff < 0.05 * moved, msg
    # inverse depths up to the common scale
    k = np.isfinite(d_o) & np.isfinite(d_r1) & (d_r1 > 1e-3)
    dd = np.abs(d_o[k] / sc - d_r1[k]) / d_r1[k]
    dd_ref = np.abs(d_r2[k] / _sim3_max_diff(c_r2, c_r1)[1] - d_r1[k]) / d_r1[k]
    assert np.median(dd) <= max(5 * np.median(dd_ref), 1e-4), (np.median(dd), np.median(dd_ref), msg)


# ------------------------------------------------------------------------------------------------- two devices in one process
def test_second_device_in_the_same_process(syn):
    """per-device state (opted-in dynamic shared memory of the tile kernels, SM count, status word, workspaces)
    used to be process-wide statics, so the first call on a SECOND device of the same process failed to launch.  Same inputs on
    cuda:0 and cuda:1 from one process: identical results (the kernels are deterministic), no error.  Needs a box with >= 2 GPUs."""
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU on this box")
    cc, cb = pkg().native_modules()
    altcorr = importlib.import_module(PKG + ".altcorr")
    w = syn.make_corr_workload(seed=1)
    s = syn.make_ba_scene(seed=2, n=40, M=96, edges="tracker")
    g = syn.make_ba_scene(seed=15, n=64, M=4, edges="global", pose_noise=0.005)
    t1 = int(s["t1"]); t0 = t1 - 10
    outs = []
    for d in ("cuda:0", "cuda:1", "cuda:0"):
        T = lambda a, dt=None: (torch.from_numpy(np.ascontiguousarray(a)).to(d) if dt is None else torch.from_numpy(np.ascontiguousarray(a)).to(d, dt))  # noqa: E731
        with torch.cuda.device(d):
            st = altcorr.corr_stacked(T(w["gmap"], torch.float16), (T(w["fmap0"], torch.float16), T(w["fmap1"], torch.float16)),
                                      T(w["coords"]), T(w["ii1"]), T(w["jj1"]), 3)
            ix, jx = cb.neighbors(T(s["kk"]), T(s["jj"]))
            po, pa = T(s["poses"].copy()), T(s["patches"].copy())
            cb.forward(po, pa, T(s["intrinsics"]), T(s["target"]), T(s["weight"]), T(s["lmbda"]), T(s["ii"]), T(s["jj"]), T(s["kk"]), t0, t1, 2)
            gp, gq = T(g["poses"].copy()), T(g["patches"].copy())
            cb.forward(gp, gq, T(g["intrinsics"]), T(g["target"]), T(g["weight"]), T(g["lmbda"]), T(g["ii"]), T(g["jj"]), T(g["kk"]), 1, 64, 2)
            lb = importlib.import_module("lietorch_backends")
            X = lb.expm(3, T(np.linspace(-0.3, 0.3, 60).reshape(10, 6)))
            torch.cuda.synchronize(d)
            assert st.device == torch.device(d) and po.device == torch.device(d)
            outs.append([t.cpu() for t in (st, ix, jx, po, pa, X)] + [gp.cpu()])
    for a, b in zip(outs[0][:6], outs[1][:6]):
        assert torch.equal(a, b)                        # deterministic kernels: bit-identical across devices
    for a, b in zip(outs[0][:6], outs[2][:6]):
        assert torch.equal(a, b)                        # and device 0 still works after device 1 was used
    assert (outs[0][6] - outs[1][6]).abs().max().item() < 2e-3   # global BA: atomics, reproducible to rounding only
