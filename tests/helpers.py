"""shared helpers for the GPU parity tests (not collected by pytest)"""
import ctypes
import importlib

import numpy as np
import torch

PKG = "wild-video-3d-reconstruction_b200"


def pkg():
    return importlib.import_module(PKG)


def dev(x, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    return t.to(dtype) if dtype is not None else t


def run_corr(f1, f2, coords, ii, jj, radius, dtype):
    cc, _ = pkg().native_modules()
    out, = cc.forward(dev(f1, dtype), dev(f2, dtype), dev(coords), dev(ii), dev(jj), radius)
    torch.cuda.synchronize()
    return out.float().cpu().numpy()


def run_ba(s, t0, t1, iters, patches_est=None, mu=2.0):
    """s: dict with numpy poses/patches/...; returns updated (poses, patches) as numpy"""
    _, cb = pkg().native_modules()
    poses, patches = dev(s["poses"].copy()), dev(s["patches"].copy())
    est = None if patches_est is None else dev(patches_est)
    cb.forward(poses, patches, dev(s["intrinsics"]), dev(s["target"]), dev(s["weight"]), dev(s["lmbda"]), dev(s["ii"]),
               dev(s["jj"]), dev(s["kk"]), int(t0), int(t1), int(iters), est, mu)
    torch.cuda.synchronize()
    return poses.cpu().numpy(), patches.cpu().numpy()


def quat_angle(qa, qb):
    """rotation angle between unit quaternions (x, y, z, w), in fp64 via the vector part of conj(a) * b (an arccos of the dot
    product quantises to 0 or 6.9e-4 rad in fp32)"""
    a = np.asarray(qa, np.float64); b = np.asarray(qb, np.float64)
    a = a / np.linalg.norm(a, axis=-1, keepdims=True); b = b / np.linalg.norm(b, axis=-1, keepdims=True)
    av, aw, bv, bw = a[..., :3], a[..., 3:], b[..., :3], b[..., 3:]
    v = aw * bv - bw * av - np.cross(av, bv)
    w = aw[..., 0] * bw[..., 0] + (av * bv).sum(-1)
    return 2 * np.arctan2(np.linalg.norm(v, axis=-1), np.abs(w))


def vp(a):
    return a.ctypes.data_as(ctypes.c_void_p)
