"""Band path of the global BA (csrc/ba_band.cu, the band layout of csrc/ba_global.cu): the reduced camera system S dX = y solved by a
block-banded Cholesky instead of the reference's dense torch::linalg::cholesky + cholesky_solve (dpvo/fastba/ba_cuda.cu:517-523).

Checked here, on the GPU, through the C ABI (wv_ba_band_solve) and through cuda_ba:
  * the solve against scipy's fp64 banded Cholesky on random SPD band systems, N = 1 ... 2047 poses, half-widths 2 ... 25, with and
    without the reference's damping, single-sided (short chains) and two-sided (long chains: both ends at once, a middle system
    in between); tolerance: fp32 backward error (stated at the assert);
  * a non-positive pivot raises the status flag; sizes beyond the maximum half-width are refused;
  * the band buffer built by ba_global_kernel holds the same numbers as the dense build, and nothing of S lies outside the band;
  * whole BA calls through the band path and through the dense path (WVB200_GLOBAL_DENSE=1) against the fp64 oracle of the
    reference's algorithm (oracle.ba_forward, fp64 mode), same tolerances as the dense-path tests in test_gpu_parity.py."""
import ctypes
import importlib

import numpy as np
import pytest
import torch

from helpers import PKG, dev, pkg, quat_angle, run_ba

pytestmark = pytest.mark.gpu

WV_STATUS_NOT_SPD = 2


@pytest.fixture(scope="module")
def syn():
    return importlib.import_module(PKG + ".synthetic")


def _lib():
    lib = pkg().capi()
    lib.wv_ba_band_floats.restype = ctypes.c_size_t
    lib.wv_ba_band_floats.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.wv_ba_band_factor_floats.restype = ctypes.c_size_t
    lib.wv_ba_band_factor_floats.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.wv_ba_band_solve.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                     ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    return lib


def _band_system(N, bs, seed):
    """random SPD block-banded system: `ab` in scipy's lower banded form (ab[k, c] = A[c + k, c]), y"""
    rng = np.random.default_rng(seed)
    n6, w = 6 * N, min(6 * (bs + 1), 6 * N)
    ab = rng.uniform(-0.5, 0.5, (w, n6))
    k = np.arange(w)[:, None]; c = np.arange(n6)[None]
    ab[((c + k) // 6 - c // 6 > bs) | (c + k >= n6)] = 0.0      # outside the block band / the matrix
    ab[0] = 0.6 * w + rng.uniform(-0.5, 0.5, n6)                 # diagonally dominant
    return ab, rng.uniform(-0.5, 0.5, n6)


def _to_band_layout(ab, ld, floats):
    """element (r, c) of the lower triangle at S[r * ld + c] (include/wvb200.h)"""
    w, n6 = ab.shape
    S = np.zeros(floats, np.float32)
    for k in range(w):
        c = np.arange(n6 - k)
        S[(c + k) * ld + c] = ab[k, : n6 - k]
    return S


def _matvec(ab, x):
    """A x for the symmetric matrix given by its lower band, fp64"""
    w, n6 = ab.shape
    out = ab[0] * x
    for k in range(1, w):
        out[k:] += ab[k, : n6 - k] * x[: n6 - k]
        out[: n6 - k] += ab[k, : n6 - k] * x[k:]
    return out


def _solve(lib, S, y, n6, bs, damp):
    V = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    dS, dy = torch.from_numpy(S).cuda(), torch.from_numpy(y.astype(np.float32)).cuda()
    F = torch.empty(lib.wv_ba_band_factor_floats(n6, bs), dtype=torch.float32, device="cuda")
    dX = torch.full((n6,), float("nan"), dtype=torch.float32, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.wv_ba_band_solve(V(dS), V(dy), n6, bs, damp, V(F), V(dX), V(status), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return rc, dX.cpu().numpy().astype(np.float64), int(status.item())


@pytest.mark.parametrize("N,bs,damp", [(1, 2, 1), (2, 2, 1), (3, 5, 0), (7, 25, 1), (26, 24, 1), (27, 25, 0), (40, 5, 1), (300, 19, 1),
                                       (301, 2, 0), (1024, 24, 1), (2047, 24, 1), (2047, 19, 0), (2047, 25, 1),
                                       # around the switch to the two-sided factorisation (N >= 2 bs + 16) and uneven halves
                                       (65, 25, 1), (66, 25, 1), (67, 25, 0), (20, 2, 1), (21, 2, 1), (101, 24, 1), (64, 24, 1)])
def test_band_solve_vs_scipy_banded_cholesky(N, bs, damp):
    from scipy.linalg import solveh_banded
    lib = _lib()
    n6, ld = 6 * N, (6 * (bs + 2) + 3) // 4 * 4
    ab, y = _band_system(N, bs, seed=N * 31 + bs)
    S = _to_band_layout(ab, ld, lib.wv_ba_band_floats(n6, bs))
    rc, x, status = _solve(lib, S, y, n6, bs, damp)
    assert rc == 0 and status == 0
    ab32 = ab.astype(np.float32).astype(np.float64)              # what the kernel was given
    if damp:                                                     # S_dd <- S_dd (1 + 1e-4) + 1  (ba_cuda.cu:517-518)
        ab32[0] = ab32[0] + (ab32[0] * 1e-4 + 1.0)
    y32 = y.astype(np.float32).astype(np.float64)
    xr = solveh_banded(ab32, y32, lower=True)
    # fp32 factorisation of a well-conditioned system: normwise backward error of a few ulp, and the solution itself to 1e-5 relative
    norm_a = np.abs(ab32).sum(0).max() * 2
    back = np.abs(_matvec(ab32, x) - y32).max() / (norm_a * np.abs(x).max() + np.abs(y32).max())
    assert back < 5e-7, back
    assert np.abs(x - xr).max() < 1e-5 * np.abs(xr).max()


def test_band_solve_flags_a_non_positive_pivot_and_refuses_bad_sizes():
    lib = _lib()
    # the bad pivot in the forward half, in the reverse half, in the middle system of the two-sided form, and in a single-sided solve
    for N, bs, where in ((40, 5, 3), (40, 5, 33), (40, 5, 19), (12, 5, 7)):
        n6, ld = 6 * N, (6 * (bs + 2) + 3) // 4 * 4
        ab, y = _band_system(N, bs, seed=3)
        ab[0, 6 * where + 2] = -50.0                             # one negative diagonal entry (stays negative after the damping)
        rc, x, status = _solve(lib, _to_band_layout(ab, ld, lib.wv_ba_band_floats(n6, bs)), y, n6, bs, 1)
        assert rc == 0 and status & WV_STATUS_NOT_SPD, (N, bs, where)
    mx = lib.wv_ba_band_max_bs()
    assert mx >= 25                                              # the tracker's PATCH_LIFETIME window (span 24) and the global pattern (19) fit
    S = _to_band_layout(*(_band_system(4, 2, 1)[0], (6 * 4 + 3) // 4 * 4, lib.wv_ba_band_floats(24, 2)))
    assert _solve(lib, S, np.zeros(24), 24, mx + 1, 1)[0] == -1  # WV_EINVAL: half-width beyond the maximum
    assert _solve(lib, S, np.zeros(24), 24, 1, 1)[0] == -1       # below the minimum
    assert _solve(lib, S, np.zeros(24), 23, 2, 1)[0] == -1       # not a multiple of 6
    assert _solve(lib, S, np.zeros(24), 0, 2, 1)[0] == 0         # nothing to solve


def _scene_args(s, t0, t1):
    D = {k: dev(s[k]) for k in ("intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")}
    poses, patches = dev(s["poses"].copy()), dev(s["patches"].copy())
    return poses, patches, D, (poses, patches, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1)


def _span_numpy(s, t0, t1):
    """widest frame span, among the free poses, of the observations of one patch -- source frame included (it gets a pose block)"""
    lo, hi = {}, {}
    for i, j, k in zip(s["ii"], s["jj"], s["kk"]):
        for f in (int(i), int(j)):
            if t0 <= f < t1:
                lo[k] = min(lo.get(k, f), f); hi[k] = max(hi.get(k, f), f)
    return max((hi[k] - lo[k] for k in lo), default=0)


SCENES = [dict(seed=21, n=60, M=6, edges="config4", loop_frac=0.0), dict(seed=15, n=48, M=6, edges="global"), dict(seed=12, n=24, M=6, edges="all"),
          dict(seed=22, n=200, M=4, edges="global")]


@pytest.mark.parametrize("kw", SCENES)
def test_band_build_holds_the_dense_system_and_nothing_lies_outside_the_band(syn, kw):
    _, cb = pkg().native_modules()
    s = syn.make_ba_scene(**kw)
    t0, t1 = 1, int(s["t1"])
    poses, patches, D, args = _scene_args(s, t0, t1)
    span = cb.global_span(poses, patches, D["ii"], D["jj"], D["kk"], t0, t1, True)
    assert span == _span_numpy(s, t0, t1)
    bs = cb.band_block(span)
    assert bs == max(span, 2)                                    # these scenes fit the band path
    n6 = 6 * (t1 - t0)
    Sd = cb.global_build(*args, 0, 1, False)
    Sb = cb.global_build_band(*args, 0, 1, bs)
    ld = (6 * (bs + 2) + 3) // 4 * 4
    dense = torch.tril(Sd[: n6 * n6].view(n6, n6)).double()
    r = torch.arange(n6, device="cuda")[:, None]; c = torch.arange(n6, device="cuda")[None]
    inband = (c <= r) & (r // 6 - c // 6 <= bs)
    scale = dense.abs().max().item()
    outside = dense[(c <= r) & ~inband]
    assert outside.numel() == 0 or outside.abs().max().item() == 0.0   # S is banded with half-width span
    got = Sb[(r * ld + c)[inband]].double()
    assert (got - dense[inband]).abs().max().item() <= 2e-6 * scale   # same sums, another order of the atomics
    nb = (n6 * (ld + 1) + 64 + 3) // 4 * 4
    assert (Sb[nb:] - Sd[n6 * n6:]).abs().max().item() <= 2e-6 * Sd[n6 * n6:].abs().max().item()
    # the shards' partial band buffers add up to the unsharded one
    parts = [cb.global_build_band(*args, q, 3, bs) for q in range(3)]
    assert (parts[0] + parts[1] + parts[2] - Sb).abs().max().item() <= 1e-5 * scale
    # both solvers return a dX with an fp32-sized BACKWARD error on the damped system (formed here in fp64 from the dense build);
    # the solutions themselves differ by the conditioning (the monocular scale is held only by the +1 damping): 5e-3 relative
    dXb = cb.global_solve_band(Sb, t1 - t0, bs).double()
    dXd = cb.global_solve(Sd.clone(), t1 - t0).double()
    cb.check_status(poses)
    A = dense + torch.tril(dense, -1).T
    A = A + torch.diag(torch.diagonal(A) * 1e-4 + 1.0)
    yv = Sd[n6 * n6:].double()
    back = lambda x: ((A @ x.reshape(-1) - yv).abs().max() / (A.abs().sum(1).max() * x.abs().max() + yv.abs().max())).item()  # noqa: E731
    assert back(dXb) < 5e-7 and back(dXd) < 5e-7, (back(dXb), back(dXd))
    assert (dXb - dXd).abs().max().item() <= 5e-3 * dXd.abs().max().item()


@pytest.mark.parametrize("kw", SCENES[:3])
@pytest.mark.parametrize("solver", ["band", "dense"])
def test_global_ba_through_band_and_dense_paths_vs_oracle(syn, kw, solver, monkeypatch):
    gp = importlib.import_module("test_gpu_parity")
    _, cb = pkg().native_modules()
    if solver == "dense":
        monkeypatch.setenv("WVB200_GLOBAL_DENSE", "1")
    s = syn.make_ba_scene(**kw)
    t0, t1 = 1, int(s["t1"])
    assert t1 - t0 > 16
    assert (cb.band_block(_span_numpy(s, t0, t1)) > 0) == (solver == "band")
    o, *tols = gp._global_tol(s, t0, t1, 2)
    poses, patches = run_ba(s, t0, t1, 2)
    gp._check_global(poses, patches, o, *tols)
    assert np.array_equal(poses[:t0], s["poses"][:t0])


def test_ba_sharded_band_route_equals_forward(syn):
    """fastba.BA_sharded (no process group) on a banded problem == cuda_ba.forward: same kernels, same order"""
    fb = importlib.import_module(PKG + ".fastba")
    s = syn.make_ba_scene(**SCENES[0])
    t1 = int(s["t1"])
    wp, wq = run_ba(s, 1, t1, 2)
    poses, patches = dev(s["poses"].copy()), dev(s["patches"].copy())
    fb.BA_sharded(poses, patches, dev(s["intrinsics"]), dev(s["target"]), dev(s["weight"]), dev(s["lmbda"]), dev(s["ii"]),
                  dev(s["jj"]), dev(s["kk"]), 1, t1, 2)
    torch.cuda.synchronize()
    gp, gq = poses.cpu().numpy(), patches.cpu().numpy()
    # float atomics order the sums differently from run to run; the scale gauge amplifies that (see _global_tol)
    assert np.abs(gp[:, :3] - wp[:, :3]).max() < 2e-3 and quat_angle(gp[:, 3:], wp[:, 3:]).max() < 2e-3
    assert np.abs(gq[:, 2] - wq[:, 2]).max() < 2e-2 * max(1.0, np.abs(wq[:, 2]).max())
