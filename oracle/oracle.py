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


# ----------------------------------------------------------------------------------------------------------
# projective ops (numpy restatement; byte/elementwise arithmetic, vectorised over edges)
#   dpvo/projective_ops.py:19-30 iproj, :33-49 proj, :53-68 transform (jacobian=False), :106-108 point_cloud, :111-121 flow_mag
#   over lietorch's SE3: dpvo/lietorch/include/se3.h:36-38 inv, :45-47 operator*, :53-56 act4; so3.h:31-37 every SO3 constructed
#   from data is normalised, :43-45 inv, :51-53 operator*, :55-60 rotation of a point.
#   keyframe distance / distance-based edges: dpvo/dpvo.py:383-408, :410-429.
# Pinned by tests/golden/pops_ref.npz = outputs of the reference's own Python + lietorch CPU ops (oracle/make_golden.py).
# ----------------------------------------------------------------------------------------------------------
def _qnorm(q):
    return q / np.sqrt((q * q).sum(-1, keepdims=True))


def _qmul(a, b):  # Hamilton product, (x, y, z, w)
    ax, ay, az, aw = np.moveaxis(a, -1, 0); bx, by, bz, bw = np.moveaxis(b, -1, 0)
    return np.stack([aw * bx + ax * bw + ay * bz - az * by, aw * by + ay * bw + az * bx - ax * bz,
                     aw * bz + az * bw + ax * by - ay * bx, aw * bw - ax * bx - ay * by - az * bz], -1)


def _qrot(q, p):  # so3.h:55-60
    v, w = q[..., :3], q[..., 3:]
    uv = np.cross(v, p)
    uv = uv + uv
    return p + w * uv + np.cross(v, uv)


def _lie_relative(Pi, Pj):
    """poses[j] * poses[i].inv() in lietorch's construction order -> (t, q)"""
    qi = _qnorm(Pi[..., 3:])
    qc = _qnorm(qi * np.array([-1, -1, -1, 1], qi.dtype))
    tinv = -_qrot(qc, Pi[..., :3])
    qc2 = _qnorm(qc)
    qj = _qnorm(Pj[..., 3:])
    q = _qnorm(_qmul(qj, qc2))
    t = Pj[..., :3] + _qrot(qj, tinv)
    return t, q


def _act_project(t, q, X0, K, depth=False):
    """act4 then proj; t, q: [E, 3|4]; X0 [E, P, P, 4]; K [E, 4]"""
    q = _qnorm(q)[:, None, None]
    X1 = _qrot(q, X0[..., :3]) + t[:, None, None] * X0[..., 3:]
    Z = X1[..., 2]
    d = 1.0 / np.maximum(Z, X0.dtype.type(0.1))
    fx, fy, cx, cy = (K[:, c, None, None] for c in range(4))
    x = fx * (d * X1[..., 0]) + cx
    y = fy * (d * X1[..., 1]) + cy
    return (np.stack([x, y, d], -1) if depth else np.stack([x, y], -1)), Z


def _pops_prep(poses, patches, intrinsics, dtype):
    P = int(np.asarray(patches.shape)[-1])
    return _np(poses, dtype).reshape(-1, 7), _np(patches, dtype).reshape(-1, 3, P, P), _np(intrinsics, dtype).reshape(-1, 4)


def _iproj(pa, K):  # pa [E, 3, P, P], K [E, 4]
    fx, fy, cx, cy = (K[:, c, None, None] for c in range(4))
    x, y, d = pa[:, 0], pa[:, 1], pa[:, 2]
    return np.stack([(x - cx) / fx, (y - cy) / fy, np.ones_like(d), d], -1)


def pops_transform(poses, patches, intrinsics, ii, jj, kk, depth=False, valid=False, tonly=False, dtype=np.float64):
    po, pa, K = _pops_prep(poses, patches, intrinsics, dtype)
    ii, jj, kk = _np(ii, np.int64), _np(jj, np.int64), _np(kk, np.int64)
    X0 = _iproj(pa[kk], K[ii])
    t, q = _lie_relative(po[ii], po[jj])
    if tonly:
        q = np.broadcast_to(np.array([0, 0, 0, 1], dtype), q.shape).copy()
    x1, Z = _act_project(t, q, X0, K[jj], depth)
    if valid:
        return x1[None], (Z > 0.2).astype(dtype)[None]
    return x1[None]


def pops_flow_mag(poses, patches, intrinsics, ii, jj, kk, beta=0.3, dtype=np.float64):
    c0 = pops_transform(poses, patches, intrinsics, ii, ii, kk, dtype=dtype)
    c1 = pops_transform(poses, patches, intrinsics, ii, jj, kk, dtype=dtype)
    c2 = pops_transform(poses, patches, intrinsics, ii, jj, kk, tonly=True, dtype=dtype)
    f1 = np.sqrt(((c1 - c0) ** 2).sum(-1)); f2 = np.sqrt(((c2 - c0) ** 2).sum(-1))
    return dtype(beta) * f1 + (dtype(1) - dtype(beta)) * f2


def pops_point_cloud(poses, patches, intrinsics, ix, dtype=np.float64):
    po, pa, K = _pops_prep(poses, patches, intrinsics, dtype)
    ix = _np(ix, np.int64)
    X0 = _iproj(pa, K[ix])
    Pf = po[ix]
    qc = _qnorm(_qnorm(Pf[:, 3:]) * np.array([-1, -1, -1, 1], dtype))
    tinv = -_qrot(qc, Pf[:, :3])
    qa = _qnorm(qc)[:, None, None]
    xyz = _qrot(qa, X0[..., :3]) + tinv[:, None, None] * X0[..., 3:]
    return np.concatenate([xyz, X0[..., 3:]], -1)[None]


def frame_distance(poses, patches, intrinsics, n, M, beta=0.5, dtype=np.float64):
    D = np.zeros((n, n), dtype)
    for i in range(n):
        for j in range(i + 1, n):
            a = np.full(M, i, np.int64); b = np.full(M, j, np.int64)
            fij = pops_flow_mag(poses, patches, intrinsics, a, b, np.arange(M * i, M * (i + 1)), beta, dtype)
            fji = pops_flow_mag(poses, patches, intrinsics, b, a, np.arange(M * j, M * (j + 1)), beta, dtype)
            D[i, j] = D[j, i] = 0.5 * (fij.mean() + fji.mean())
    return D


def distance_based_edges(D, thresh):
    n = D.shape[0]
    ei, ej = list(range(n - 1)), list(range(1, n))
    for i in range(n):
        for j in range(i + 2, n):
            if D[i, j] < thresh:
                ei.append(i); ej.append(j)
    return np.array(ei, np.int64), np.array(ej, np.int64)


# ------------------------------------------------------------------------------------------------- SoftAgg (dpvo/blocks.py:31-48)
# The reference's arithmetic lives in a dependency that is NOT in /root/reference: torch-scatter==2.1.2 (environment.yaml:43).
# Its published algorithm (torch_scatter/composite/softmax.py::scatter_softmax at that tag) is restated below:
#     max_value_per_index = scatter_max(src, index)[0];  recentered = src - max_value_per_index.gather(index)
#     e = recentered.exp();  sum_per_index = scatter_sum(e, index);  return e / sum_per_index.gather(index)
# followed by blocks.py:43  y = scatter_sum(f(x) * w, jx, dim=1).  PARITY UNPINNED for the softmax/sum arithmetic (torch_scatter
# cannot be imported here); the grouping itself (`torch.unique(ix, return_inverse=True)`, blocks.py:41) is pinned against
# torch.unique in tests/test_oracle_golden.py.
def unique_inverse(ix):
    """blocks.py:41: `_, jx = torch.unique(ix, return_inverse=True)` -> (sorted distinct keys, jx)"""
    u, jx = np.unique(np.asarray(ix, np.int64), return_inverse=True)
    return u, jx.astype(np.int64).reshape(-1)


def _seq_scatter(acc, jx, src):
    """acc[:, jx[e]] += src[:, e] for e ascending, rounding to acc.dtype after every addition"""
    for e in range(src.shape[1]):
        acc[:, jx[e]] = (acc[:, jx[e]] + src[:, e]).astype(acc.dtype)
    return acc


def soft_agg(fx, gx, ix, dtype=np.float64):
    """fx, gx [B, E, D]; returns (y [B, G, D], jx).  dtype=np.float16 rounds after every step like the reference's fp16 tensors
    do under autocast (scatter_sum accumulating in ascending edge order); np.float64 is the exact value of the same formula."""
    _, jx = unique_inverse(ix)
    G = int(jx.max()) + 1 if jx.size else 0
    f = np.asarray(fx).astype(dtype); g = np.asarray(gx).astype(dtype)
    B, E, D = g.shape
    mx = np.full((B, G, D), -np.inf, dtype)
    for e in range(E):
        mx[:, jx[e]] = np.maximum(mx[:, jx[e]], g[:, e])
    ex = np.exp((g - mx[:, jx]).astype(dtype)).astype(dtype)
    den = _seq_scatter(np.zeros((B, G, D), dtype), jx, ex)
    w = (ex / den[:, jx]).astype(dtype)
    y = _seq_scatter(np.zeros((B, G, D), dtype), jx, (f * w).astype(dtype))
    return y, jx
