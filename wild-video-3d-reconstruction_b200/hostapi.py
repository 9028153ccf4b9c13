"""Host-buffer front end of the C ABI (include/wvb200.h: wv_session_*), for embedders without torch tensors.

    s = hostapi.Session()
    s.update(gmap, [fmap_l0, fmap_l1], coords, ii1, jj1, corr_out=[out0, out1], ba=dict(poses=..., patches=..., ...))
    s.update(gmap, [fmap_l0, fmap_l1], coords, ii1, jj1, None, corr_stacked_out=out882, ba=...)     # DPVO.corr's stacked layout

One call = one tracker update() on the replaced path (dpvo/dpvo.py:711-749): altcorr.corr on every pyramid level,
fastba.neighbors and fastba.BA on the same edge list.  All arrays are HOST arrays (numpy, or pinned torch CPU tensors);
`poses` / `patches` are updated in place like fastba.BA does.  There is no CPU fallback: the session needs libwvb200.so
and an sm_100 device.
"""
import ctypes

import numpy as np

from . import capi

WV_F16, WV_F32 = 0, 1


class UpdateArgs(ctypes.Structure):
    _fields_ = [
        ("dtype", ctypes.c_int), ("B", ctypes.c_int), ("E", ctypes.c_int), ("C", ctypes.c_int), ("P", ctypes.c_int),
        ("radius", ctypes.c_int), ("n1", ctypes.c_int), ("n2", ctypes.c_int), ("nlevels", ctypes.c_int),
        ("fmap1", ctypes.c_void_p), ("fmap2", ctypes.c_void_p * 4), ("H2", ctypes.c_int * 4), ("W2", ctypes.c_int * 4),
        ("coord_scale", ctypes.c_float * 4), ("coords", ctypes.c_void_p), ("ii1", ctypes.c_void_p), ("jj1", ctypes.c_void_p),
        ("corr_out", ctypes.c_void_p * 4),
        ("poses", ctypes.c_void_p), ("patches", ctypes.c_void_p), ("intrinsics", ctypes.c_void_p), ("target", ctypes.c_void_p),
        ("weight", ctypes.c_void_p), ("lmbda", ctypes.c_void_p), ("ii", ctypes.c_void_p), ("jj", ctypes.c_void_p),
        ("kk", ctypes.c_void_p), ("n_poses", ctypes.c_int64), ("n_patches", ctypes.c_int64), ("t0", ctypes.c_int),
        ("t1", ctypes.c_int), ("iterations", ctypes.c_int), ("patches_est", ctypes.c_void_p), ("mu", ctypes.c_float),
        ("nb_ix", ctypes.c_void_p), ("nb_jx", ctypes.c_void_p), ("status", ctypes.c_int),
        ("corr_stacked_out", ctypes.c_void_p),
    ]


def _ptr(a):
    if a is None:
        return None
    if hasattr(a, "data_ptr"):  # torch CPU tensor (pinned or not)
        assert not a.is_cuda and a.is_contiguous()
        return ctypes.c_void_p(a.data_ptr())
    assert a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


def _is_half(a):
    return str(a.dtype) in ("float16", "torch.float16")


class Session:
    def __init__(self):
        self.lib = capi()
        self.lib.wv_session_create.restype = ctypes.c_void_p
        self.lib.wv_session_destroy.argtypes = [ctypes.c_void_p]
        self.lib.wv_session_update_host.argtypes = [ctypes.c_void_p, ctypes.POINTER(UpdateArgs)]
        self.h = self.lib.wv_session_create()
        if not self.h:
            raise RuntimeError("wv_session_create failed (no CUDA device?)")

    def close(self):
        if self.h:
            self.lib.wv_session_destroy(self.h)
            self.h = None

    __del__ = close

    def make_args(self, fmap1, fmap2_levels, coords, ii1, jj1, corr_out, radius=3, coord_scale=None, ba=None, iterations=2,
                  corr_stacked_out=None):
        """corr_out: one [B,E,7,7,P,P] array per level -- or None with corr_stacked_out = ONE [B,E,7*7*P*P*2] array in DPVO.corr's
        stacked layout (two levels, tracker configuration)"""
        a = UpdateArgs()
        B, n1, C, P, _ = fmap1.shape
        a.dtype = WV_F16 if _is_half(fmap1) else WV_F32
        a.B, a.E, a.C, a.P, a.radius, a.n1 = B, coords.shape[1], C, P, radius, n1
        a.n2 = fmap2_levels[0].shape[1]
        a.nlevels = len(fmap2_levels)
        a.fmap1 = _ptr(fmap1)
        for l, f in enumerate(fmap2_levels):
            a.fmap2[l] = _ptr(f).value
            a.H2[l], a.W2[l] = f.shape[3], f.shape[4]
            a.coord_scale[l] = (coord_scale[l] if coord_scale is not None else 0.25 ** l)
            if corr_out is not None:
                a.corr_out[l] = _ptr(corr_out[l]).value
        a.corr_stacked_out = _ptr(corr_stacked_out)
        assert corr_out is not None or corr_stacked_out is not None
        a.coords, a.ii1, a.jj1 = _ptr(coords), _ptr(ii1), _ptr(jj1)
        a.iterations = 0
        if ba is not None:
            for k in ("poses", "patches", "intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk"):
                setattr(a, k, _ptr(ba[k]))
            a.n_poses = int(np.prod(ba["poses"].shape)) // 7
            a.n_patches = int(np.prod(ba["patches"].shape)) // (3 * P * P)
            a.t0, a.t1, a.iterations = int(ba["t0"]), int(ba["t1"]), int(iterations)
            a.patches_est = _ptr(ba.get("patches_est"))
            a.mu = float(ba.get("mu", 2.0))
            a.nb_ix, a.nb_jx = _ptr(ba.get("nb_ix")), _ptr(ba.get("nb_jx"))
        return a

    def run(self, args):
        rc = self.lib.wv_session_update_host(self.h, ctypes.byref(args))
        if rc != 0:
            self.lib.wv_strerror.restype = ctypes.c_char_p
            raise RuntimeError(f"wv_session_update_host: {self.lib.wv_strerror(rc).decode()} (status {args.status})")
        if args.status & 2:
            raise RuntimeError("linalg.cholesky: the reduced camera system is not positive-definite")
        return args.status

    def update(self, fmap1, fmap2_levels, coords, ii1, jj1, corr_out, **kw):
        return self.run(self.make_args(fmap1, fmap2_levels, coords, ii1, jj1, corr_out, **kw))
