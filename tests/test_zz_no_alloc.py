"""wv_neighbors_static (include/wvb200.h): the no-allocation variant of cuda_ba.neighbors -- scratch in a static device array inside
libwvb200.so -- called through the C ABI and compared with the existing entry point wv_neighbors (caller-supplied workspace).
Integer work: the outputs must be bit-identical.  Two calls in a row on DIFFERENT edge lists (the second one smaller and with other
ids), so that stale contents of the shared scratch would show.  Collected last on purpose."""
import ctypes
import importlib

import numpy as np
import pytest
import torch

PKG = "wild-video-3d-reconstruction_b200"
MAX_EDGES, MAX_GROUPS = 262144, 1048576   # WV_NEIGHBORS_STATIC_MAX_EDGES / _GROUPS
WV_ELIMIT = -5


def _lib():
    lib = importlib.import_module(PKG).capi()
    i64p, vp = ctypes.c_void_p, ctypes.c_void_p
    lib.wv_neighbors.argtypes = [i64p, i64p, ctypes.c_int64, ctypes.c_int64, i64p, i64p, vp, ctypes.c_size_t, vp, vp]
    lib.wv_neighbors_static.argtypes = [i64p, i64p, ctypes.c_int64, ctypes.c_int64, i64p, i64p, vp, vp]
    lib.wv_ba_workspace_bytes.restype = ctypes.c_size_t
    lib.wv_ba_workspace_bytes.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.c_int]
    return lib


def _edges(seed, n_frames, M, lifetime, removal, offset):
    """tracker-shaped edge list (patch-major blocks, shuffled), patch ids shifted by `offset`"""
    syn = importlib.import_module(PKG + ".synthetic")
    s = syn.make_ba_scene(seed=seed, n=n_frames, M=M, edges="tracker", lifetime=lifetime, removal=removal)
    rng = np.random.default_rng(seed)
    perm = rng.permutation(len(s["kk"]))
    return (s["kk"][perm] + offset).astype(np.int64), s["jj"][perm].astype(np.int64)


def _both(lib, kk, jj, n_groups):
    V = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    E = len(kk)
    dkk, djj = torch.from_numpy(kk).cuda(), torch.from_numpy(jj).cuda()
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # existing entry point: workspace from the caller
    wsb = lib.wv_ba_workspace_bytes(E, n_groups, 0)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    ix0, jx0 = torch.empty(E, dtype=torch.int64, device="cuda"), torch.empty(E, dtype=torch.int64, device="cuda")
    assert lib.wv_neighbors(V(dkk), V(djj), E, n_groups, V(ix0), V(jx0), V(ws), wsb, V(status), st) == 0
    # new entry point: no workspace, nothing allocated inside (outputs were allocated above, before the call)
    ix1 = torch.full((E,), -7, dtype=torch.int64, device="cuda")
    jx1 = torch.full((E,), -7, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()
    assert lib.wv_neighbors_static(V(dkk), V(djj), E, n_groups, V(ix1), V(jx1), V(status), st) == 0
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == before
    assert int(status.item()) == 0
    return (ix0.cpu().numpy(), jx0.cpu().numpy()), (ix1.cpu().numpy(), jx1.cpu().numpy()), free_before - free_after


@pytest.mark.gpu
def test_neighbors_static_matches_workspace_entry_point_twice_in_a_row():
    import oracle as orc
    lib = _lib()
    # call 1: default.yaml-sized list (M = 384: E = 181 248), ids up to 15 360; call 2: a smaller list with other ids
    cases = [_edges(11, 40, 384, 13, 22, 0) + (1 << 16,), _edges(12, 14, 8, 6, 9, 5000) + (8192,)]
    assert len(cases[0][0]) == 181248 and len(cases[0][0]) <= MAX_EDGES
    grown = []
    for kk, jj, n_groups in cases:
        ref, new, used = _both(lib, kk, jj, n_groups)
        assert np.array_equal(ref[0], new[0]) and np.array_equal(ref[1], new[1])    # bit-exact, like every index result
        grown.append(used)
    # the second list against the CPU oracle too (ba.cpp:106-151 restated), so that both entry points cannot be wrong together
    oi, oj = orc.neighbors(cases[1][0], cases[1][1])
    assert np.array_equal(new[0], oi) and np.array_equal(new[1], oj)
    # device memory did not shrink across the second call (the first may have triggered lazy module loading)
    assert grown[1] < (1 << 20)


@pytest.mark.gpu
def test_neighbors_static_rejects_sizes_beyond_its_scratch():
    lib = _lib()
    lib.wv_strerror.restype = ctypes.c_char_p
    d = torch.zeros(8, dtype=torch.int64, device="cuda")
    V = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    assert lib.wv_neighbors_static(V(d), V(d), MAX_EDGES + 1, 1024, V(d), V(d), None, None) == WV_ELIMIT
    assert lib.wv_neighbors_static(V(d), V(d), 8, MAX_GROUPS + 1, V(d), V(d), None, None) == WV_ELIMIT
    assert b"fixed maximum" in lib.wv_strerror(WV_ELIMIT)
    assert lib.wv_neighbors_static(V(d), V(d), 0, 1024, V(d), V(d), None, None) == 0      # empty list: nothing to do
    assert lib.wv_neighbors_static(None, V(d), 8, 1024, V(d), V(d), None, None) == -1     # WV_EINVAL
