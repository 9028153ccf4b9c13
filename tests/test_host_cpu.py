"""CPU-only tests of the host logic and of the C ABI surface (no compute calls: there is no GPU here)."""
import ctypes
import importlib
import os
import re
import socket

import numpy as np
import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "wild-video-3d-reconstruction_b200"


def _pkg():
    return importlib.import_module(PKG)


def test_c_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "wvb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(wv_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 18, names
    lib = _pkg().capi()
    for n in names:
        assert hasattr(lib, n), f"libwvb200.so does not export {n}"


def test_c_abi_argument_validation_without_a_device():
    lib = _pkg().capi()
    lib.wv_strerror.restype = ctypes.c_char_p
    assert lib.wv_strerror(0) == b"ok" and b"invalid" in lib.wv_strerror(-1)
    assert lib.wv_device_supported() in (0, 1)
    # empty problems are OK before any pointer is looked at; nonsense sizes are rejected
    assert lib.wv_corr_forward(None, None, None, None, None, None, 0, 1, 0, 128, 3, 3, 1, 1, 8, 8, 3, None, ctypes.c_size_t(0), None, None) == 0
    assert lib.wv_corr_forward(None, None, None, None, None, None, 7, 1, 4, 128, 3, 3, 1, 1, 8, 8, 3, None, ctypes.c_size_t(0), None, None) == -1
    assert lib.wv_corr_forward(None, None, None, None, None, None, 0, 1, 4, 128, 3, 3, 1, 1, 8, 8, 3, None, ctypes.c_size_t(0), None, None) == -1
    lib.wv_ba_workspace_bytes.restype = ctypes.c_size_t
    lib.wv_ba_workspace_bytes.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.c_int]
    assert lib.wv_ba_workspace_bytes(45312, 3840, 10) > 45312 * 8
    lib.wv_corr_workspace_bytes.restype = ctypes.c_size_t
    assert lib.wv_corr_workspace_bytes(1, 1, 45312, 128, 3, 3, 3456, 36, 96, 128, 3) == 0  # fp32 maps: generic kernel, no staging


@pytest.mark.parametrize("tag,freen", [("full", -1), ("free4", 4)])
def test_solve_system_host_entry_matches_reference(golden, tag, freen):
    """cuda_ba.solve_system is host code in the reference too (ba.cpp:156-234): the C ABI entry runs here"""
    lib = _pkg().capi()
    g = golden["ba_ref"]
    Ji, Jj = np.ascontiguousarray(g["pgo_Ji"], np.float32), np.ascontiguousarray(g["pgo_Jj"], np.float32)
    ii, jj = np.ascontiguousarray(g["pgo_ii"], np.int64), np.ascontiguousarray(g["pgo_jj"], np.int64)
    res = np.ascontiguousarray(g["pgo_res"], np.float32)
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
    f = lib.wv_solve_system_host
    f.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int64, ctypes.c_float, ctypes.c_float, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    n = ctypes.c_int64(0)
    assert f(vp(Ji), vp(Jj), vp(ii), vp(jj), vp(res), len(ii), 1e-4, 1e-4, freen, None, ctypes.byref(n)) == 0
    delta = np.zeros((n.value, 7), np.float32)
    assert f(vp(Ji), vp(Jj), vp(ii), vp(jj), vp(res), len(ii), 1e-4, 1e-4, freen, vp(delta), ctypes.byref(n)) == 0
    assert np.abs(delta - g[f"pgo_{tag}_delta"]).max() < 1e-5
    bad = ii.copy(); bad[0] = jj[0]
    assert f(vp(Ji), vp(Jj), vp(bad), vp(jj), vp(res), len(ii), 1e-4, 1e-4, freen, None, ctypes.byref(n)) == -1  # i == j


def test_tracker_edge_list_and_neighbors_invariants(syn):
    """bit-exact structure: the synthetic edge list follows dpvo.py:756-769 / :657-658; neighbors chains are consistent"""
    ii, jj, kk = syn.tracker_edges(40, 96)
    assert len(kk) == 45312 and (ii == kk // 96).all()
    assert ((kk // 96) >= 40 - 22).all() and (np.abs(jj - kk // 96) <= 12).all()
    ix, jx = oracle.neighbors(kk, jj)
    has_next = jx >= 0
    assert (ix[jx[has_next]] == np.nonzero(has_next)[0]).all()          # next's previous is me
    assert (kk[jx[has_next]] == kk[has_next]).all()                     # same patch
    assert (jj[jx[has_next]] >= jj[has_next]).all()                     # frames in order
    assert (ix < 0).sum() == len(np.unique(kk)) == (jx < 0).sum()       # one head and one tail per patch


def test_clip_sharding():
    d = _pkg().dist
    assert d.clip_shard(64, 3, 8) == list(range(24, 32))
    allc = sum((d.clip_shard(10, r, 4) for r in range(4)), [])
    assert allc == list(range(10))


def _gloo_worker(rank, world, port, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import sys
    sys.path.insert(0, ROOT)
    d = importlib.import_module(PKG).dist
    r, w, _ = d.init("gloo")
    t = d.max_over_ranks(1.0 + r)            # bench.py's time reduction
    d.barrier()
    bench = importlib.import_module("bench")
    wk, s = bench.make_inputs(r, small=True)  # rank r tracks its own clip
    q.put((r, w, t, float(wk["coords"].sum()), d.clip_shard(5, r, w)))
    d.finalize()


def test_two_ranks_gloo_replicas():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    out = sorted(q.get(timeout=120) for _ in ps)
    [p.join(60) for p in ps]
    assert [o[0] for o in out] == [0, 1] and all(o[1] == 2 for o in out)
    assert all(o[2] == 2.0 for o in out)                    # max over ranks
    assert out[0][3] != out[1][3]                           # different clips per rank
    assert out[0][4] + out[1][4] == list(range(5))          # shards partition the clips


# ----------------------------------------------------------------------------------------------------------
# fastba.BA_sharded host logic under gloo (world_size 2): the collective placement, shard arguments and iteration loop, with a
# CPU stand-in for the four cuda_ba.global_* pieces (a linear toy problem with the same build -> solve -> dz -> retract algebra).
# ----------------------------------------------------------------------------------------------------------
class _ToyGlobalBA:
    def __init__(self, band=0):
        self.calls = []
        self.band = band                        # what band_block answers: 0 = dense route, > 0 = band route

    def global_span(self, poses, patches, ii, jj, kk, t0, t1, build_plan):
        self.calls.append(("span", bool(build_plan)))
        return 5

    def band_block(self, span):
        assert span == 5
        return self.band

    def global_build_band(self, poses, patches, intr, target, weight, lmbda, ii, jj, kk, t0, t1, shard, nshards, bs, est, mu):
        # the toy problem is dense; its "band buffer" is the dense one -- what is under test is which calls BA_sharded makes
        assert bs == self.band
        self.calls.append(("build_band", shard, nshards, bs))
        self.calls, keep = [], self.calls
        out = self.global_build(poses, patches, intr, target, weight, lmbda, ii, jj, kk, t0, t1, shard, nshards, False, est, mu)
        self.calls = keep
        return out

    def global_solve_band(self, Sy, n_free, bs):
        assert bs == self.band
        self.calls.append(("solve_band", bs))
        return self.global_solve(Sy, n_free)

    @staticmethod
    def _patch_terms(k, depth, n6):
        import torch
        g = torch.Generator().manual_seed(int(k) + 17)
        e = torch.randn(n6, generator=g, dtype=torch.float64) * 0.3
        b = torch.randn(n6, generator=g, dtype=torch.float64) * 0.1
        C = 1.0 + 0.1 * float(depth)           # depends on the current state: iteration 2 sees iteration 1's update
        u = 0.5 * float(depth) - 0.2
        return e, b, C, u

    def _shard(self, kk, shard, nshards):
        import torch
        kx = torch.unique(kk)                   # sorted
        lo, hi = len(kx) * shard // nshards, len(kx) * (shard + 1) // nshards
        return kx, lo, hi

    def global_build(self, poses, patches, intr, target, weight, lmbda, ii, jj, kk, t0, t1, shard, nshards, build_plan, est, mu):
        import torch
        self.calls.append(("build", shard, nshards, bool(build_plan)))
        n6 = 6 * (t1 - t0)
        S = torch.zeros(n6, n6, dtype=torch.float64); y = torch.zeros(n6, dtype=torch.float64)
        kx, lo, hi = self._shard(kk, shard, nshards)
        for r in range(lo, hi):
            e, b, C, u = self._patch_terms(kx[r], patches[kx[r], 2, 0, 0], n6)
            q = 1.0 / (C + float(lmbda[0]))
            S += torch.outer(b, b) + torch.outer(e, e) * (1.0 - q * 0.5)
            y += b * 0.3 - q * u * e
        return torch.cat([torch.tril(S).reshape(-1), y]).float()

    def global_solve(self, Sy, n_free):
        import torch
        n6 = 6 * n_free
        S = Sy[: n6 * n6].view(n6, n6).double(); y = Sy[n6 * n6:].double()
        S = torch.tril(S) + torch.tril(S, -1).T
        S = S + torch.diag(torch.diagonal(S) * 1e-4 + 1.0)
        return torch.cholesky_solve(y[:, None], torch.linalg.cholesky(S)).view(n_free, 6).float()

    def global_dz(self, poses, patches, intr, target, weight, lmbda, ii, jj, kk, t0, t1, shard, nshards, dX, est, mu):
        import torch
        self.calls.append(("dz", shard, nshards))
        n6 = 6 * (t1 - t0)
        kx, lo, hi = self._shard(kk, shard, nshards)
        dZ = torch.zeros(min(ii.numel(), patches.shape[0]))
        for r in range(lo, hi):
            e, b, C, u = self._patch_terms(kx[r], patches[kx[r], 2, 0, 0], n6)
            dZ[r] = float((u - (e * dX.double().view(-1)).sum()) / (C + float(lmbda[0])))
        return dZ

    def check_status(self, like):
        self.calls.append(("check",))

    def global_retract(self, poses, patches, E, t0, t1, dX, dZ):
        import torch
        kx = self._kx
        poses[t0:t1, :6] += dX
        patches[kx, 2] += dZ[: len(kx), None, None]


def _sharded_worker(rank, world, port, q, band=0):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import sys
    sys.path.insert(0, ROOT)
    import torch
    pkg = importlib.import_module(PKG)
    if world > 1:
        pkg.dist.init("gloo")
    ba = importlib.import_module(PKG + ".fastba.ba")
    toy = _ToyGlobalBA(band)
    ba.cuda_ba = toy                                   # the CPU stand-in; the product path is the CUDA extension
    n, M = 9, 5
    ii, jj, kk = pkg.synthetic.global_edges(n, M)
    ii, jj, kk = (torch.from_numpy(x) for x in (ii, jj, kk))
    toy._kx = torch.unique(kk)
    g = torch.Generator().manual_seed(3)
    poses = torch.randn(n, 7, generator=g); patches = torch.rand(n * M, 3, 3, 3, generator=g) + 0.5
    E = ii.numel()
    ba.BA_sharded(poses, patches, torch.ones(n, 4), torch.zeros(E, 2), torch.ones(E, 2), torch.tensor([1e-4]), ii, jj, kk, 1, n, 2)
    q.put((rank, poses.numpy().copy(), patches.numpy().copy(), toy.calls))
    if world > 1:
        pkg.dist.finalize()


@pytest.mark.parametrize("band", [0, 3])
def test_ba_sharded_host_logic_gloo_two_ranks(band):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    results = {}
    for world in (1, 2):
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
        q = ctx.Queue()
        ps = [ctx.Process(target=_sharded_worker, args=(r, world, port, q, band)) for r in range(world)]
        [p.start() for p in ps]
        results[world] = sorted((q.get(timeout=180) for _ in ps), key=lambda o: o[0])
        [p.join(60) for p in ps]
    one, (r0, r1) = results[1][0], results[2]
    # both ranks end with the same state, bit for bit (same reduced system after the allreduce, same solve, same retraction)
    assert np.array_equal(r0[1], r1[1]) and np.array_equal(r0[2], r1[2])
    # and it is the one-process answer up to the summation order of the allreduce
    np.testing.assert_allclose(r0[1], one[1], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(r0[2], one[2], rtol=1e-5, atol=1e-6)
    assert not np.array_equal(one[1], results[1][0][1] * 0)  # something moved
    # shard arguments and the plan flag: (rank, world) on every call; the span query builds the plan once, the builds reuse it;
    # the band route is taken iff band_block says so, with the matching solve
    for rank, res in enumerate((r0, r1)):
        assert [c for c in res[3] if c[0] == "span"] == [("span", True)]
        if band:
            assert [c for c in res[3] if c[0] == "build_band"] == [("build_band", rank, 2, band)] * 2
            assert [c for c in res[3] if c[0] == "solve_band"] == [("solve_band", band)] * 2
            assert not [c for c in res[3] if c[0] == "build"]
        else:
            builds = [c for c in res[3] if c[0] == "build"]
            assert [b[1:3] for b in builds] == [(rank, 2), (rank, 2)] and [b[3] for b in builds] == [False, False]
            assert not [c for c in res[3] if c[0] in ("build_band", "solve_band")]
        assert [c[1:] for c in res[3] if c[0] == "dz"] == [(rank, 2), (rank, 2)]


@pytest.mark.parametrize("n,loops,freen", [(2048, 24, -1), (700, 5, 500), (40, 0, -1)])
def test_solve_system_at_pose_graph_scale_vs_sparse_fp64(n, loops, freen, ref_modules):
    """a pose graph of n frames: the chain plus `loops` long-range edges (one of them 0 <-> n - 1), the size at which the dense
    (7n)^2 matrix of round 1 would have been 1.6 GB.  The entry point keeps the system in skyline storage; checked against scipy's sparse
    fp64 solve of the same normal equations and -- where oracle/_ref exists -- against the reference's own Eigen-based solve_system"""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    rng = np.random.default_rng(n)
    ii = np.arange(n - 1, dtype=np.int64); jj = ii + 1
    if loops:
        li = rng.integers(0, n // 2, loops); lj = li + rng.integers(n // 4, n // 2, loops)
        li[0], lj[0] = 0, n - 1
        ii, jj = np.concatenate([ii, li]), np.concatenate([jj, lj])
        flip = rng.random(len(ii)) < 0.3                      # edge direction is arbitrary in the reference
        ii, jj = np.where(flip, jj, ii).astype(np.int64), np.where(flip, ii, jj).astype(np.int64)
    r = len(ii)
    Ji = (np.eye(7)[None] + 0.1 * rng.standard_normal((r, 7, 7))).astype(np.float32)
    Jj = (-np.eye(7)[None] + 0.1 * rng.standard_normal((r, 7, 7))).astype(np.float32)
    res = rng.standard_normal((r, 7)).astype(np.float32)
    ep, lm = 1e-4, 1e-4
    lib = _pkg().capi()
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
    f = lib.wv_solve_system_host
    f.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int64, ctypes.c_float, ctypes.c_float, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    nn = ctypes.c_int64(0)
    delta = np.zeros((n, 7), np.float32)
    assert f(vp(Ji), vp(Jj), vp(ii), vp(jj), vp(res), r, ep, lm, freen, vp(delta), ctypes.byref(nn)) == 0 and nn.value == n
    # the same normal equations in scipy: J is [7r x 7n] with two 7x7 blocks per edge row (ba.cpp:174-234)
    rows = (np.arange(r)[:, None, None] * 7 + np.arange(7)[None, :, None]) + np.zeros((1, 1, 7), np.int64)
    cols_i = ii[:, None, None] * 7 + np.arange(7)[None, None, :] + np.zeros((1, 7, 1), np.int64)
    cols_j = jj[:, None, None] * 7 + np.arange(7)[None, None, :] + np.zeros((1, 7, 1), np.int64)
    J = sp.coo_matrix((np.concatenate([Ji.astype(np.float64).ravel(), Jj.astype(np.float64).ravel()]),
                       (np.concatenate([rows.ravel(), rows.ravel()]), np.concatenate([cols_i.ravel(), cols_j.ravel()]))), shape=(7 * r, 7 * n)).tocsr()
    A = (J.T @ J).tolil()
    d = A.diagonal()
    A.setdiag(d + d * np.float64(np.float32(lm)) + np.float64(np.float32(ep)))
    b = -(J.T @ res.astype(np.float64).ravel())
    m = 7 * n if freen < 0 else 7 * freen
    x = np.zeros(7 * n)
    x[:m] = spl.spsolve(A.tocsc()[:m, :m], b[:m])
    assert np.abs(delta.ravel() - x).max() < 2e-5 * max(1.0, np.abs(x).max())
    assert np.all(delta.ravel()[m:] == 0)
    ref = ref_modules.get("ref_cuda_ba") if hasattr(ref_modules, "get") else None
    if ref is not None and n <= 700:                          # the reference itself (Eigen SimplicialLLT), where it was built
        import torch
        dr, = ref.solve_system(torch.from_numpy(Ji), torch.from_numpy(Jj), torch.from_numpy(ii), torch.from_numpy(jj), torch.from_numpy(res), ep, lm, freen)
        assert np.abs(delta.ravel() - dr.numpy().ravel()).max() < 2e-5 * max(1.0, np.abs(x).max())


# ------------------------------------------------------------------------------------------------- COLMAP text export
def test_float_formatter_is_cpython_repr():
    """every number of the export goes through wv_format_f64; the reference formats with f-strings / str(): CPython's float repr"""
    import random
    import struct
    lib = ctypes.CDLL(os.path.join(ROOT, PKG, "libwvb200.so"))
    lib.wv_format_f64.argtypes = [ctypes.c_double, ctypes.c_char_p]
    buf = ctypes.create_string_buffer(64)
    rnd = random.Random(7)
    vals = [0.0, -0.0, 1.0, -1.5, 1e16, 1e15, 9999999999999998.0, 1e17, 1e-4, 1e-5, 0.0001234, 123456.789, 5e-324, 1.7976931348623157e308,
            float("inf"), float("-inf"), float("nan"), 0.1, 1 / 3, 2.5e-7, 1e22, 1e21, 123456789012345678.0]
    vals += [struct.unpack("d", struct.pack("Q", rnd.getrandbits(64)))[0] for _ in range(20000)]
    vals += [float(np.float32(rnd.uniform(-100, 100))) for _ in range(20000)]
    vals += [rnd.uniform(-1, 1) * 10 ** rnd.randint(-8, 20) for _ in range(20000)]
    for v in vals:
        n = lib.wv_format_f64(v, buf)
        assert buf.raw[:n].decode() == repr(v), v
    for v in vals[:5000]:   # the pose columns are numpy float64 scalars in an f-string: same text
        assert f"{np.float64(v)}" == repr(v)


@pytest.mark.parametrize("case", ["plain", "named"])
def test_colmap_export_is_byte_identical_to_the_reference(case, tmp_path):
    """tests/golden/colmap_ref/*.npz: inputs + the three files written by the reference's own save_output_for_COLMAP
    (oracle/make_golden_colmap.py)"""
    ce = importlib.import_module(PKG + ".colmap_export")
    g = np.load(os.path.join(ROOT, "tests", "golden", "colmap_ref", case + ".npz"))
    poses, tstamps = g["poses"], g["tstamps"]
    if bool(g["with_images"]):
        os.makedirs(tmp_path / "images")
        for i in range(len(poses)):
            open(tmp_path / "images" / f"frame_{i:05d}.{'png' if i % 2 else 'jpg'}", "w").close()
    traj = ce.Trajectory(positions_xyz=poses[:, :3], orientations_quat_wxyz=poses[:, [6, 3, 4, 5]], timestamps=tstamps)   # dpvo_demo.py:187-192
    ce.save_output_for_COLMAP(str(tmp_path / "colmap"), tstamps, traj, g["points"], g["colors"], False, *g["calib"], 384, 512)
    for f in ("images.txt", "points3D.txt", "cameras.txt"):
        want = bytes(g[f.replace(".", "_")])
        got = open(tmp_path / "colmap" / f, "rb").read()
        assert got == want, f
    with pytest.raises(NotImplementedError):
        ce.save_output_for_COLMAP(str(tmp_path / "colmap"), tstamps, traj, g["points"], g["colors"], True, *g["calib"], 384, 512)


def test_colmap_export_large_and_edge_cases(tmp_path):
    ce = importlib.import_module(PKG + ".colmap_export")
    rng = np.random.default_rng(0)
    n = 200000
    pts = (rng.standard_normal((n, 3)) * 50).astype(np.float32)
    col = rng.integers(0, 256, (n, 3)).astype(np.uint8)
    txt = ce.points3d_text(pts, col)
    lines = txt.split("\n")
    assert len(lines) == n + 1 and lines[-1] == ""
    for i in (0, 1, 777, n - 1):    # the reference's row expression, plot_utils.py:91-92
        p, c = pts[i].tolist(), col[i].tolist()
        assert lines[i] == f"{i + 1} " + " ".join(map(str, p + c)) + " 0.0 0 0 0 0 0 0"
    assert ce.points3d_text(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.uint8)) == ""
    assert ce.images_text(np.zeros((0, 4)), np.zeros((0, 3)), []) == ""
    q = rng.standard_normal(4); q /= np.linalg.norm(q)
    M = ce.quaternion_matrix(q)
    q2 = ce.quaternion_from_matrix(M)
    assert np.allclose(q2, q if q[0] >= 0 else -q, atol=1e-12)
