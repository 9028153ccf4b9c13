"""Seeded synthetic workloads for the altcorr + fastba hot path (numpy only).

Shapes follow SURVEY.md section 8(d): 512x384 frames -> 1/4-resolution maps of 96x128 (level 0)
and 24x32 (level 1), 128 channels, 3x3 patches, radius 3; edge lists follow the tracker's own
rule (dpvo/dpvo.py:756-769 __edges_forw/__edges_back, removal at :657-658).
Used by tests/, bench.py and oracle/make_golden.py.  No oracle or reference import here.
"""
import numpy as np


# ----------------------------------------------------------------------------- SE3 helpers (float64)
def quat_mul(a, b):
    ax, ay, az, aw = np.moveaxis(a, -1, 0)
    bx, by, bz, bw = np.moveaxis(b, -1, 0)
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz], -1)


def quat_rot(q, v):
    qv = q[..., :3]
    uv = 2.0 * np.cross(qv, v)
    return v + q[..., 3:4] * uv + np.cross(qv, uv)


def so3_exp(phi):
    th = np.linalg.norm(phi, axis=-1, keepdims=True)
    small = th < 1e-8
    ths = np.where(small, 1.0, th)
    imag = np.where(small, 0.5, np.sin(0.5 * ths) / ths)
    return np.concatenate([imag * phi, np.cos(0.5 * th)], -1)


def se3_mul(a, b):
    """(ta,qa)*(tb,qb): x -> Ra(Rb x + tb) + ta; pose layout [tx,ty,tz,qx,qy,qz,qw]"""
    t = quat_rot(a[..., 3:], b[..., :3]) + a[..., :3]
    return np.concatenate([t, quat_mul(a[..., 3:], b[..., 3:])], -1)


def se3_inv(a):
    qi = a[..., 3:] * np.array([-1, -1, -1, 1.0])
    return np.concatenate([-quat_rot(qi, a[..., :3]), qi], -1)


def se3_act(a, p):
    return quat_rot(a[..., 3:], p) + a[..., :3]


# ----------------------------------------------------------------------------- edge lists
def tracker_edges(n, M, lifetime=13, removal=22):
    """Edge list the unchanged tracker holds after its n-th frame (no keyframe culling).

    Replays append_factors(__edges_forw) + append_factors(__edges_back) for every frame and the
    removal rule `ii < n - REMOVAL_WINDOW`.  Patch k belongs to frame k // M (index_ is the identity
    without culling).  Returns int64 (ii, jj, kk) in the tracker's own order.
    """
    ii_l, jj_l, kk_l = [], [], []
    for cur in range(1, n + 1):  # cur == self.n after the increment at dpvo.py:852
        r = lifetime
        # __edges_forw: patches of frames [cur-r, cur-1) -> frame cur-1  (indexing='ij': patch-major)
        k0, k1 = M * max(cur - r, 0), M * max(cur - 1, 0)
        kf = np.arange(k0, k1, dtype=np.int64)
        kk_l.append(kf); jj_l.append(np.full_like(kf, cur - 1))
        # __edges_back: patches of frame cur-1 -> frames [cur-r, cur)
        kb = np.arange(M * max(cur - 1, 0), M * cur, dtype=np.int64)
        fr = np.arange(max(cur - r, 0), cur, dtype=np.int64)
        kk_l.append(np.repeat(kb, len(fr))); jj_l.append(np.tile(fr, len(kb)))
        kk = np.concatenate(kk_l); jj = np.concatenate(jj_l)
        keep = (kk // M) >= cur - removal
        kk_l, jj_l = [kk[keep]], [jj[keep]]
    kk, jj = kk_l[0], jj_l[0]
    return kk // M, jj, kk


def all_pairs_edges(n, M):
    """every patch -> every frame of the window incl. its own (BASELINE.json config 1: E = n*n*M)"""
    kk = np.repeat(np.arange(n * M, dtype=np.int64), n)
    jj = np.tile(np.arange(n, dtype=np.int64), n * M)
    return kk // M, jj, kk


def global_edges(n, M):
    """Edge list of the reference's global BA in its fixed-pattern form (dpvo/dpvo.py:448-471): frame pairs (i, i+1) for all i,
    plus (i, j) for every 5th i and j in [i+10, i+20); each pair carries all M patches of frame i, pair-major."""
    fi = [np.arange(n - 1, dtype=np.int64)]
    fj = [np.arange(1, n, dtype=np.int64)]
    for i in range(0, n, 5):
        j = np.arange(i + 10, min(i + 20, n), dtype=np.int64)
        fi.append(np.full_like(j, i)); fj.append(j)
    fi, fj = np.concatenate(fi), np.concatenate(fj)
    ii = np.repeat(fi, M)
    jj = np.repeat(fj, M)
    kk = (fi[:, None] * M + np.arange(M, dtype=np.int64)[None]).reshape(-1)
    return ii, jj, kk


def config4_edges(n, M, span=12, loop_frac=0.02, seed=3):
    """every patch -> the frames within +-span of its source frame (25 per patch away from the ends),
    plus loop_frac extra edges to frames drawn uniformly from [0, n).  Patch-major."""
    rng = np.random.default_rng(seed)
    src = np.repeat(np.arange(n, dtype=np.int64), M)                      # source frame of every patch
    off = np.arange(-span, span + 1, dtype=np.int64)
    jj = src[:, None] + off[None]
    kk = np.broadcast_to(np.arange(n * M, dtype=np.int64)[:, None], jj.shape)
    ok = (jj >= 0) & (jj < n)
    jj, kk = jj[ok], kk[ok]
    nl = int(loop_frac * len(kk))
    kl = rng.integers(0, n * M, nl).astype(np.int64)
    jl = rng.integers(0, n, nl).astype(np.int64)
    kk = np.concatenate([kk, kl]); jj = np.concatenate([jj, jl])
    return kk // M, jj, kk


# ----------------------------------------------------------------------------- BA scene
def make_ba_scene(seed=0, n=10, M=96, edges="all", ht=96, wd=128, fx=110.0, n_buffer=None,
                  pose_noise=0.02, depth_noise=0.1, pix_noise=0.5, first_frame=0, lifetime=13, removal=22, loop_frac=0.02,
                  heavy_patches=()):
    """Synthetic window: forward motion 0.1 m/frame + 1 deg/frame yaw, points at depth U(2,8) m.

    Returns a dict of float32/int64 numpy arrays with the layouts cuda_ba.forward takes:
    poses [Nbuf,7] (world->camera), patches [Nbuf*M,3,3,3], intrinsics [Nbuf,4], target/weight [E,2],
    ii/jj/kk [E], plus poses_gt / patches_gt / patches_est (prior with 20 % invalid = 0).
    Frames live at buffer rows first_frame .. first_frame+n-1.  `heavy_patches`: patch ids (0-based, before `first_frame`) that
    EVERY other frame observes, whatever the edge pattern says (a camera turning on the spot under distance-based edges).
    """
    rng = np.random.default_rng(seed)
    P = 3
    nb = n_buffer or (first_frame + n)
    cx, cy = wd / 2.0, ht / 2.0
    # ground-truth camera-to-world trajectory, then invert (poses are world->camera)
    t = np.arange(n)[:, None]
    yaw = np.deg2rad(1.0) * t
    c2w = np.concatenate([0.02 * np.sin(0.7 * t), 0.01 * np.cos(0.5 * t), 0.1 * t,
                          so3_exp(np.concatenate([0 * yaw, yaw, 0 * yaw], -1))], -1)
    w2c = se3_inv(c2w)
    # patches: centre uniform in [1, wd-2] x [1, ht-2], 3x3 grid of unit offsets, true depth U(2,8)
    ctr = np.stack([rng.uniform(1, wd - 2, n * M), rng.uniform(1, ht - 2, n * M)], -1)
    depth = rng.uniform(2.0, 8.0, n * M)
    gy, gx = np.meshgrid(np.arange(-1, 2), np.arange(-1, 2), indexing="ij")
    patches = np.zeros((nb * M, 3, P, P))
    rows = first_frame * M + np.arange(n * M)
    patches[rows, 0] = ctr[:, 0, None, None] + gx
    patches[rows, 1] = ctr[:, 1, None, None] + gy
    patches[rows, 2] = (1.0 / depth)[:, None, None]
    patches_gt = patches.copy()
    if edges == "all":
        ii, jj, kk = all_pairs_edges(n, M)
    elif edges == "global":
        ii, jj, kk = global_edges(n, M)
    elif edges == "config4":
        ii, jj, kk = config4_edges(n, M, loop_frac=loop_frac)
    else:
        ii, jj, kk = tracker_edges(n, M, lifetime, removal)
    for k in heavy_patches:
        keep = kk != k
        oth = np.setdiff1d(np.arange(n), [k // M])
        ii = np.concatenate([ii[keep], np.full(len(oth), k // M)]); jj = np.concatenate([jj[keep], oth]); kk = np.concatenate([kk[keep], np.full(len(oth), k)])
    # true projections of the patch centres
    src = kk // M
    Xc = np.stack([(ctr[kk, 0] - cx) / fx, (ctr[kk, 1] - cy) / fx, np.ones(len(kk))], -1) * depth[kk, None]
    Xw = se3_act(c2w[src], Xc)
    Xj = se3_act(w2c[jj], Xw)
    proj = np.stack([fx * Xj[:, 0] / Xj[:, 2] + cx, fx * Xj[:, 1] / Xj[:, 2] + cy], -1)
    target = proj + rng.normal(0, pix_noise, proj.shape)
    weight = rng.uniform(0, 1, proj.shape)
    # perturbed initial state
    poses = np.zeros((nb, 7)); poses[:, 6] = 1.0
    poses_gt = poses.copy()
    poses_gt[first_frame:first_frame + n] = w2c
    xi = rng.normal(0, pose_noise, (n, 6))
    dT = np.concatenate([xi[:, :3], so3_exp(xi[:, 3:])], -1)
    poses[first_frame:first_frame + n] = se3_mul(dT, w2c)
    poses[first_frame] = w2c[0]
    patches[rows, 2] *= (1.0 + rng.normal(0, depth_noise, n * M))[:, None, None]
    est = patches_gt.copy()
    est[rows, 2] *= (1.0 + rng.normal(0, 0.05, n * M))[:, None, None]
    est[rows[rng.uniform(size=n * M) < 0.2], 2] = 0.0
    intr = np.tile(np.array([fx, fx, cx, cy]), (nb, 1))
    f32 = np.float32
    return dict(poses=poses.astype(f32), patches=patches.astype(f32), intrinsics=intr.astype(f32),
                target=target.astype(f32), weight=weight.astype(f32), lmbda=np.array([1e-4], f32),
                ii=ii + first_frame, jj=jj + first_frame, kk=kk + first_frame * M,
                poses_gt=poses_gt.astype(f32), patches_gt=patches_gt.astype(f32), patches_est=est.astype(f32),
                t0=first_frame + 1, t1=first_frame + n, M=M, n=n)


# ----------------------------------------------------------------------------- corr workload
def make_corr_workload(seed=1, M=96, pmem=36, n_frames=40, ht=96, wd=128, C=128, E_limit=None,
                       oob_frac=0.05, lifetime=13, removal=22):
    """BASELINE.json config 2: gmap/fmap ring in the tracker's layout and the tracker's edge list.

    Returns float32 arrays (the caller casts maps to fp16): gmap [1,pmem*M,C,3,3], fmap0 [1,pmem,C,ht,wd],
    fmap1 = 4x4 average pool [1,pmem,C,ht/4,wd/4], coords [1,E,2,3,3] at level-0 resolution (level 1 uses
    coords/4), and ring indices ii1 = kk % (M*pmem), jj1 = jj % pmem exactly as dpvo/dpvo.py:326-333.
    """
    rng = np.random.default_rng(seed)
    f32 = np.float32
    gmap = (0.25 * rng.standard_normal((1, pmem * M, C, 3, 3), dtype=f32))
    fmap0 = (0.25 * rng.standard_normal((1, pmem, C, ht, wd), dtype=f32))
    fmap1 = fmap0.reshape(1, pmem, C, ht // 4, 4, wd // 4, 4).mean(axis=(4, 6)).astype(f32)
    ii, jj, kk = tracker_edges(n_frames, M, lifetime, removal)
    if E_limit is not None:
        ii, jj, kk = ii[:E_limit], jj[:E_limit], kk[:E_limit]
    E = len(kk)
    ctr = np.stack([rng.uniform(1, wd - 2, n_frames * M), rng.uniform(1, ht - 2, n_frames * M)], -1)
    flow = rng.normal(0, 2.0, (E, 2)) + rng.uniform(0, 1, (E, 2))
    c = ctr[kk] + flow
    far = rng.uniform(size=E) < oob_frac  # fully / partly out of bounds
    c[far] += rng.choice([-1.0, 1.0], (far.sum(), 2)) * rng.uniform(0.5 * wd, 1.2 * wd, (far.sum(), 2))
    gy, gx = np.meshgrid(np.arange(-1, 2), np.arange(-1, 2), indexing="ij")
    coords = np.stack([c[:, 0, None, None] + gx, c[:, 1, None, None] + gy], 1)[None].astype(f32)
    return dict(gmap=gmap, fmap0=fmap0, fmap1=fmap1, coords=coords,
                ii1=(kk % (M * pmem)).astype(np.int64), jj1=(jj % pmem).astype(np.int64),
                ii=ii, jj=jj, kk=kk, E=E, M=M, pmem=pmem)
