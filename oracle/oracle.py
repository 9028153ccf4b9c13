"""TEST INFRASTRUCTURE ONLY -- ctypes front end of the CPU oracle (oracle/liboracle.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does.  See wv_oracle.h for the citations of
the reference lines each function follows and for how the oracle is pinned.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
F64, F32, F16EMU = 0, 1, 2
_lib = None


def build(force=False):
    so = os.path.join(HERE, "liboracle.so")
    srcs = [os.path.join(HERE, f) for f in ("oracle_corr.c", "oracle_ba.c", "oracle_ba_impl.inc", "wv_oracle.h")]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _np(x, dtype):
    if hasattr(x, "detach"):
        x = x.detach().cpu()
        if str(x.dtype) in ("torch.float16", "torch.bfloat16"):
            x = x.float()
        x = x.numpy()
    return np.ascontiguousarray(x, dtype=dtype)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def corr_forward(fmap1, fmap2, coords, ii, jj, radius, mode=F64, nthreads=0):
    """-> [B, E, 2r+1, 2r+1, PH, PW] (dim2 = x-tap, dim3 = y-tap); float64 for F64 else float32."""
    f1, f2, co = _np(fmap1, np.float32), _np(fmap2, np.float32), _np(coords, np.float32)
    ii, jj = _np(ii, np.int64), _np(jj, np.int64)
    B, n1, C, PH, PW = f1.shape
    _, n2, _, H2, W2 = f2.shape
    E = co.shape[1]
    T = 2 * radius + 1
    out = np.empty((B, E, T, T, PH, PW), dtype=np.float64 if mode == F64 else np.float32)
    rc = lib().orc_corr_forward(_p(f1), _p(f2), _p(co), _p(ii), _p(jj), _p(out), B, E, C, PH, PW, n1, n2,
                                H2, W2, radius, mode, nthreads)
    if rc:
        raise IndexError("corr_forward: edge index out of range")
    return out


def corr_backward(fmap1, fmap2, coords, ii, jj, grad, radius):
    f1, f2, co = _np(fmap1, np.float32), _np(fmap2, np.float32), _np(coords, np.float32)
    ii, jj, g = _np(ii, np.int64), _np(jj, np.int64), _np(grad, np.float32)
    B, n1, C, PH, PW = f1.shape
    _, n2, _, H2, W2 = f2.shape
    E = co.shape[1]
    g1, g2 = np.empty(f1.shape, np.float64), np.empty(f2.shape, np.float64)
    rc = lib().orc_corr_backward(_p(f1), _p(f2), _p(co), _p(ii), _p(jj), _p(g), _p(g1), _p(g2), B, E, C, PH,
                                 PW, n1, n2, H2, W2, radius)
    if rc:
        raise IndexError("corr_backward: edge index out of range")
    return g1, g2


def patchify_forward(net, coords, radius):
    net, co = _np(net, np.float32), _np(coords, np.float32)
    B, C, H, W = net.shape
    M = co.shape[1]
    D = 2 * radius + 2
    out = np.empty((B, M, C, D, D), np.float32)
    lib().orc_patchify_forward(_p(net), _p(co), _p(out), B, M, C, H, W, radius)
    return out


def patchify_backward(net_shape, coords, gradient, radius):
    co, g = _np(coords, np.float32), _np(gradient, np.float32)
    B, C, H, W = net_shape
    M = co.shape[1]
    out = np.empty((B, C, H, W), np.float64)
    lib().orc_patchify_backward(_p(g), _p(co), _p(out), B, M, C, H, W, radius)
    return out


def ba_forward(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, iterations=2,
               patches_est=None, mu=2.0, mode=F64):
    """Functional form: returns dict(poses, patches, dX, dZ, Mp, status); inputs are not modified."""
    po = _np(poses, np.float32).reshape(-1, 7).copy()
    P = int(np.asarray(patches.shape)[-1])
    pa = _np(patches, np.float32).reshape(-1, 3, P, P).copy()
    it = _np(intrinsics, np.float32).reshape(-1, 4)
    tg, wt = _np(target, np.float32).reshape(-1, 2), _np(weight, np.float32).reshape(-1, 2)
    lm = _np(lmbda, np.float32).reshape(-1)
    ii, jj, kk = _np(ii, np.int64), _np(jj, np.int64), _np(kk, np.int64)
    E = ii.shape[0]
    pe = None if patches_est is None else _np(patches_est, np.float32).reshape(-1, 3, P, P)
    N = t1 - t0
    Mp_max = len(np.unique(kk))
    dX = np.zeros((iterations, max(N, 0), 6), np.float64)
    dZ = np.zeros((iterations, Mp_max), np.float64)
    Mp = ctypes.c_int64(0)
    f = lib().orc_ba_forward
    f.argtypes = [ctypes.c_void_p] * 9 + [ctypes.c_int64] * 3 + [ctypes.c_int] * 4 + \
                 [ctypes.c_void_p, ctypes.c_float, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    rc = f(_p(po), _p(pa), _p(it), _p(tg), _p(wt), _p(lm), _p(ii), _p(jj), _p(kk), E, po.shape[0], pa.shape[0],
           P, t0, t1, iterations, _p(pe), float(mu), mode, _p(dX), _p(dZ), ctypes.byref(Mp))
    if rc < 0:
        raise IndexError("ba_forward: index out of range")
    return dict(poses=po, patches=pa, dX=dX, dZ=dZ, Mp=Mp.value, status=rc)


def reproject(poses, patches, intrinsics, ii, jj, kk, mode=F64):
    po = _np(poses, np.float32).reshape(-1, 7)
    P = int(np.asarray(patches.shape)[-1])
    pa = _np(patches, np.float32).reshape(-1, 3, P, P)
    it = _np(intrinsics, np.float32).reshape(-1, 4)
    ii, jj, kk = _np(ii, np.int64), _np(jj, np.int64), _np(kk, np.int64)
    E = ii.shape[0]
    out = np.empty((1, E, 2, P, P), np.float32)
    f = lib().orc_reproject
    f.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int64] * 3 + [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    if f(_p(po), _p(pa), _p(it), _p(ii), _p(jj), _p(kk), E, po.shape[0], pa.shape[0], P, _p(out), mode):
        raise IndexError("reproject: index out of range")
    return out


def neighbors(ii, jj):
    ii, jj = _np(ii, np.int64), _np(jj, np.int64)
    E = ii.shape[0]
    ix, jx = np.empty(E, np.int64), np.empty(E, np.int64)
    f = lib().orc_neighbors
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]
    f(_p(ii), _p(jj), E, _p(ix), _p(jx))
    return ix, jx


def solve_system(J_i, J_j, ii, jj, res, ep, lm, freen):
    Ji, Jj = _np(J_i, np.float32).reshape(-1, 7, 7), _np(J_j, np.float32).reshape(-1, 7, 7)
    ii, jj, rs = _np(ii, np.int64), _np(jj, np.int64), _np(res, np.float32).reshape(-1, 7)
    r = rs.shape[0]
    n = int(max(ii.max(), jj.max())) + 1
    delta = np.zeros((n, 7), np.float32)
    nout = ctypes.c_int64(0)
    f = lib().orc_solve_system
    f.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int64, ctypes.c_float, ctypes.c_float, ctypes.c_int,
                                           ctypes.c_void_p, ctypes.c_void_p]
    rc = f(_p(Ji), _p(Jj), _p(ii), _p(jj), _p(rs), r, float(ep), float(lm), int(freen), _p(delta), ctypes.byref(nout))
    if rc < 0:
        raise ValueError("solve_system: i == j or negative index")
    return delta, rc
