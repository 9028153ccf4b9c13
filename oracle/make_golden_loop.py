#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY -- freezes the REFERENCE's output for the loop-closure call shape of fastba.BA as
tests/golden/ba_loop_ref.npz.

dpvo/loop_closure/long_term.py:104-127 calls fastba.BA in a shape no tracker call has: 3 poses (frames i-1, i, i+1, all FIXED:
t0 = t1 = 3, the structure-only branch), n keypoints of the middle frame as 3x3 "patches" that all start at ONE inverse depth
(the median of the frame), two edges per patch with ii == 1 and jj in {0, 2}, kk = arange(n).repeat(2) -- i.e. the edge list is
NOT grouped by patch --, unit weights, lmbda = 1e-3, 6 iterations, intrinsics x4 (full-resolution pixels).

The inputs are built here exactly like that (same tensor expressions); the outputs come from the reference's own ba_cuda.cu / ba.cpp
compiled as host code (oracle/_ref/ref_cuda_ba.so, build_ref.py).  Runs only in the authoring container."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import build_ref  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def qmul(a, b):
    ax, ay, az, aw = a; bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz])


def qrot(q, v):
    x, y, z, w = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    return v @ R.T


def scene(seed, n, outliers):
    """three world->camera poses around a forward-moving, slightly yawing camera; n points seen from the middle frame"""
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = 4 * 110.0, 4 * 112.0, 4 * 64.0, 4 * 48.0      # the tracker's 1/4-resolution intrinsics x 4 (long_term.py:118)
    poses = np.zeros((3, 7), np.float32)
    for f in range(3):
        yaw = np.deg2rad(2.0 * (f - 1)) + 0.01 * rng.standard_normal()
        q = np.array([0.0, np.sin(yaw / 2), 0.0, np.cos(yaw / 2)])
        q = qmul(q, np.array([0.01 * rng.standard_normal(), 0.0, 0.0, 1.0])); q /= np.linalg.norm(q)
        c = np.array([0.15 * (f - 1) + 0.02 * rng.standard_normal(), 0.01 * rng.standard_normal(), 0.35 * (f - 1)])
        poses[f, :3] = -qrot(q, c); poses[f, 3:] = q              # t = -R c
    u = rng.uniform(20, 4 * 128 - 20, n); v = rng.uniform(20, 4 * 96 - 20, n)
    depth = rng.uniform(2.0, 9.0, n)
    Xc = np.stack([(u - cx) / fx * depth, (v - cy) / fy * depth, depth], 1)          # in the middle camera
    q1, t1 = poses[1, 3:].astype(np.float64), poses[1, :3].astype(np.float64)
    Xw = qrot(np.array([-q1[0], -q1[1], -q1[2], q1[3]]), Xc - t1)
    kps = []
    for f in (0, 2):
        q, t = poses[f, 3:].astype(np.float64), poses[f, :3].astype(np.float64)
        Xf = qrot(q, Xw) + t
        kps.append(np.stack([fx * Xf[:, 0] / Xf[:, 2] + cx, fy * Xf[:, 1] / Xf[:, 2] + cy], 1) + 0.4 * rng.standard_normal((n, 2)))
    bad = rng.choice(n, outliers, replace=False)                   # wrong matches: the call's residual filter exists for these
    kps[0][bad] += rng.uniform(-40, 40, (outliers, 2))
    true_disp = float(np.median(1.0 / depth))
    # the tensor expressions of long_term.py:104-116
    kk = torch.arange(n).repeat(2)
    ii = torch.ones(2 * n, dtype=torch.long)
    jj = torch.zeros(2 * n, dtype=torch.long); jj[n:] = 2
    kps1 = torch.from_numpy(np.stack([u, v], 1).astype(np.float32))[None]
    patches = torch.cat((kps1, torch.ones(1, n, 1) * true_disp), dim=-1)
    patches = patches[:, :, :, None, None].repeat(1, 1, 1, 3, 3)                     # '1 n uvd -> 1 n uvd 3 3'
    target = torch.stack((torch.from_numpy(kps[0].astype(np.float32))[None], torch.from_numpy(kps[1].astype(np.float32))[None]))
    target = target.permute(1, 0, 2, 3).reshape(1, 2 * n, 2)                           # 'ot 1 n uv -> 1 (ot n) uv'
    weight = torch.ones_like(target)
    intr = torch.tensor([[fx, fy, cx, cy]], dtype=torch.float32).repeat(3, 1)
    return dict(poses=poses, patches=patches[0].numpy().copy(), intrinsics=intr.numpy(), target=target[0].numpy().copy(),
                weight=weight[0].numpy().copy(), lmbda=np.array([1e-3], np.float32), ii=ii.numpy(), jj=jj.numpy(), kk=kk.numpy(),
                depth_gt=(1.0 / depth).astype(np.float32), bad=bad.astype(np.int64))


def main():
    build_ref.build()
    ba = build_ref.load("ref_cuda_ba")
    T = torch.from_numpy
    out = {}
    for name, (seed, n, outliers) in {"lc257": (1, 257, 9), "lc40": (2, 40, 0), "lc1": (3, 1, 0)}.items():
        s = scene(seed, n, outliers)
        poses, patches = T(s["poses"].copy()), T(s["patches"].copy())
        ba.forward(poses, patches, T(s["intrinsics"]), T(s["target"]), T(s["weight"]), T(s["lmbda"]), T(s["ii"]), T(s["jj"]), T(s["kk"]), 3, 3, 6)
        assert np.array_equal(poses.numpy(), s["poses"])            # structure only: poses untouched
        out.update({f"{name}_in_{k}": v for k, v in s.items()})
        out.update({f"{name}_poses": poses.numpy(), f"{name}_patches": patches.numpy()})
        d = patches.numpy()[:, 2, 0, 0]
        good = np.setdiff1d(np.arange(n), s["bad"])
        print(name, "n", n, "median |d - d_gt| / d_gt before", float(np.median(np.abs(s["patches"][good, 2, 0, 0] - s["depth_gt"][good]) / s["depth_gt"][good])),
              "after", float(np.median(np.abs(d[good] - s["depth_gt"][good]) / s["depth_gt"][good])))
    np.savez_compressed(os.path.join(OUT, "ba_loop_ref.npz"), **out)
    print("ba_loop_ref.npz", os.path.getsize(os.path.join(OUT, "ba_loop_ref.npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
