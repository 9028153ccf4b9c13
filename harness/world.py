"""Synthetic 512x384 clip for BASELINE.json configs[2..4]: analytic ray casting, no graphics API.

A textured room (the inside of an axis-aligned box) with free-standing axis-aligned pillars, seen from a camera that walks a
smooth closed 6-DoF path at ~0.25 m per frame (5 FPS walking speed).  Every pixel's depth is the exact ray/box intersection, the
texture is a procedural 3-D value noise evaluated at the hit point (so it is view-consistent by construction), and 1-3 rectangles
drift over the image as "dynamic objects": they are painted with their own texture, get a fake near depth, and are marked invalid
in the mask (mask semantics: non-zero = static / valid, dpvo/stream.py:67-70, dpvo/net.py:141).  Depth is clamped at 10x its median
like dpvo/stream.py:63-64.

Everything is torch (runs on the CPU in the tests and on the GPU in the bench); geometry is fp64, images uint8 BGR [H, W, 3] as
cv2.imread would give, depth fp32 [H, W], mask bool [H, W], intrinsics fp64 [4] -- the tuple dpvo_demo.py:107-110 feeds DPVO.

Conventions: camera x right, y down, z forward; poses are world->camera (X_c = R X_w + t) as (tx, ty, tz, qx, qy, qz, qw), the
layout of PatchGraph.poses_ (dpvo/patchgraph.py:28).
"""
import math

import numpy as np
import torch


def _hash3(ix, iy, iz, seed):
    h = ix * 374761393 + iy * 668265263 + iz * 2147483647 + seed * 1274126177
    h = (h ^ (h >> 13)) * 1274126177
    h = h ^ (h >> 16)
    return (h & 0xFFFFFF).to(torch.float32) / 16777216.0


def value_noise3(p, seed, octaves=4, freq=1.3):
    """p [..., 3] (fp32/fp64) -> [...] in [0, 1]: sum of trilinear lattice noise over `octaves` octaves"""
    out = torch.zeros(p.shape[:-1], dtype=torch.float32, device=p.device)
    amp, tot = 1.0, 0.0
    for o in range(octaves):
        q = p * (freq * 2.0 ** o)
        f = torch.floor(q)
        w = (q - f).to(torch.float32)
        w = w * w * (3.0 - 2.0 * w)
        i = f.to(torch.int64)
        acc = 0.0
        for dx in (0, 1):
            for dy in (0, 1):
                for dz in (0, 1):
                    h = _hash3(i[..., 0] + dx, i[..., 1] + dy, i[..., 2] + dz, seed + 101 * o)
                    wx = w[..., 0] if dx else 1.0 - w[..., 0]
                    wy = w[..., 1] if dy else 1.0 - w[..., 1]
                    wz = w[..., 2] if dz else 1.0 - w[..., 2]
                    acc = acc + h * wx * wy * wz
        out = out + amp * acc
        tot += amp
        amp *= 0.55
    return out / tot


def quat_from_R(R):
    """[..., 3, 3] rotation -> (x, y, z, w), fp64 numpy"""
    R = np.asarray(R, np.float64)
    q = np.empty(R.shape[:-2] + (4,))
    tr = R[..., 0, 0] + R[..., 1, 1] + R[..., 2, 2]
    for idx in np.ndindex(*R.shape[:-2]):
        m = R[idx]
        t = tr[idx]
        if t > 0:
            s = math.sqrt(t + 1.0) * 2
            w, x, y, z = 0.25 * s, (m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s
        elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
            s = math.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
            w, x, y, z = (m[2, 1] - m[1, 2]) / s, 0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s
        elif m[1, 1] > m[2, 2]:
            s = math.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
            w, x, y, z = (m[0, 2] - m[2, 0]) / s, (m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s
        else:
            s = math.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
            w, x, y, z = (m[1, 0] - m[0, 1]) / s, (m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s
        q[idx] = (x, y, z, w)
    return q


class World:
    def __init__(self, seed=2, ht=384, wd=512, n_frames=1000, device="cpu", speed=0.25, n_dynamic=2, n_pillars=7):
        self.seed, self.ht, self.wd, self.n_frames = int(seed), int(ht), int(wd), int(n_frames)
        self.device = torch.device(device)
        rng = np.random.default_rng(seed)
        f = 0.86 * wd
        self.intrinsics = np.array([f, f, 0.5 * wd, 0.5 * ht], np.float64)
        # room half extents (x, y, z) and the walking ellipse
        self.room = np.array([15.0, 3.2, 11.0])
        self.a, self.b = 9.5 + rng.uniform(-0.5, 0.5), 6.0 + rng.uniform(-0.5, 0.5)
        # pillars: axis-aligned boxes standing on the floor (y = +room_y is the floor: y points down)
        boxes = []
        tries = 0
        while len(boxes) < n_pillars and tries < 1000:
            tries += 1
            cx, cz = rng.uniform(-13, 13), rng.uniform(-9, 9)
            r = math.hypot(cx / self.a, cz / self.b)
            if abs(r - 1.0) < 0.28:       # keep the walking path clear
                continue
            hw, hd, hh = rng.uniform(0.4, 1.1), rng.uniform(0.4, 1.1), rng.uniform(1.0, 3.0)
            lo = np.array([cx - hw, self.room[1] - 2 * hh, cz - hd])
            hi = np.array([cx + hw, self.room[1], cz + hd])
            boxes.append((lo, hi))
        self.box_lo = torch.tensor(np.array([b[0] for b in boxes]), dtype=torch.float64, device=self.device)
        self.box_hi = torch.tensor(np.array([b[1] for b in boxes]), dtype=torch.float64, device=self.device)
        # trajectory
        k = np.arange(n_frames, dtype=np.float64)
        mean_r = 0.5 * (self.a + self.b)
        s = 0.35 + k * speed / mean_r
        pos = np.stack([self.a * np.cos(s), 0.6 + 0.35 * np.sin(0.017 * k * speed / 0.25), self.b * np.sin(s)], -1)
        tang = np.stack([-self.a * np.sin(s), np.zeros_like(s), self.b * np.cos(s)], -1)
        tang /= np.linalg.norm(tang, axis=-1, keepdims=True)
        yaw = 0.45 + 0.18 * np.sin(0.05 * k * speed / 0.25)        # look a little outward of the tangent
        pitch = 0.07 * np.sin(0.031 * k * speed / 0.25 + 1.0)
        roll = 0.05 * np.sin(0.023 * k * speed / 0.25 + 2.0)
        R_wc = np.empty((n_frames, 3, 3))
        for i in range(n_frames):
            fwd0 = tang[i]
            c, s_ = math.cos(-yaw[i]), math.sin(-yaw[i])          # rotate about world y (down)
            fwd = np.array([c * fwd0[0] + s_ * fwd0[2], 0.0, -s_ * fwd0[0] + c * fwd0[2]])
            fwd = fwd * math.cos(pitch[i]) + np.array([0.0, 1.0, 0.0]) * math.sin(pitch[i])
            fwd /= np.linalg.norm(fwd)
            down0 = np.array([0.0, 1.0, 0.0])
            right = np.cross(down0, fwd); right /= np.linalg.norm(right)
            down = np.cross(fwd, right)
            cr, sr = math.cos(roll[i]), math.sin(roll[i])
            right, down = cr * right + sr * down, -sr * right + cr * down
            R_wc[i] = np.stack([right, down, fwd], -1)              # columns: camera axes in world coordinates
        self.R_cw = np.transpose(R_wc, (0, 2, 1)).copy()
        self.t_cw = -np.einsum("nij,nj->ni", self.R_cw, pos)
        self.pos = pos
        self.poses_cw = np.concatenate([self.t_cw, quat_from_R(self.R_cw)], -1)       # [n, 7] world->camera
        self.R_cw_t = torch.tensor(self.R_cw, dtype=torch.float64, device=self.device)
        self.t_cw_t = torch.tensor(self.t_cw, dtype=torch.float64, device=self.device)
        self.pos_t = torch.tensor(pos, dtype=torch.float64, device=self.device)
        # dynamic rectangles (image space): centre path, half size, fake depth
        self.dyn = []
        for d in range(n_dynamic):
            self.dyn.append(dict(c0=np.array([rng.uniform(0.25, 0.75) * wd, rng.uniform(0.3, 0.7) * ht]),
                                 amp=np.array([rng.uniform(0.1, 0.3) * wd, rng.uniform(0.05, 0.2) * ht]),
                                 om=rng.uniform(0.02, 0.06, 2), ph=rng.uniform(0, 6.28, 2),
                                 half=np.array([rng.uniform(0.06, 0.11) * wd, rng.uniform(0.08, 0.16) * ht]),
                                 depth=rng.uniform(1.2, 2.0), seed=int(rng.integers(1, 1 << 20))))
        yy, xx = torch.meshgrid(torch.arange(ht, dtype=torch.float64, device=self.device),
                                torch.arange(wd, dtype=torch.float64, device=self.device), indexing="ij")
        self._px, self._py = xx, yy

    # ------------------------------------------------------------------------------------------------------------------
    def gt_poses(self):
        """[n_frames, 7] world->camera (tx, ty, tz, qx, qy, qz, qw), fp64 numpy"""
        return self.poses_cw.copy()

    def gt_cam_positions(self):
        return self.pos.copy()

    def dyn_rects(self, t):
        """[(x0, y0, x1, y1, depth, seed)] of the dynamic rectangles in frame t (full-resolution pixels)"""
        out = []
        for d in self.dyn:
            c = d["c0"] + d["amp"] * np.sin(d["om"] * t + d["ph"])
            out.append((c[0] - d["half"][0], c[1] - d["half"][1], c[0] + d["half"][0], c[1] + d["half"][1], d["depth"], d["seed"]))
        return out

    def _cast(self, o, d):
        """o [..., 3] origins, d [..., 3] directions (any length) -> ray parameter of the first static hit"""
        inv = 1.0 / torch.where(d.abs() < 1e-12, torch.full_like(d, 1e-12), d)
        room = torch.tensor(self.room, dtype=torch.float64, device=d.device)
        t1 = (-room - o) * inv
        t2 = (room - o) * inv
        t_room = torch.maximum(t1, t2).min(dim=-1).values                 # exit parameter of the room (origin is inside)
        best = t_room
        lo, hi = self.box_lo.to(d.device), self.box_hi.to(d.device)
        if lo.shape[0]:                                                    # all pillars at once: [..., B, 3]
            a1 = (lo - o[..., None, :]) * inv[..., None, :]
            a2 = (hi - o[..., None, :]) * inv[..., None, :]
            tn = torch.minimum(a1, a2).max(dim=-1).values
            tf = torch.maximum(a1, a2).min(dim=-1).values
            tn = torch.where((tn < tf) & (tn > 1e-9), tn, torch.full_like(tn, float("inf")))
            best = torch.minimum(best, tn.min(dim=-1).values)
        return best

    def raycast_world(self, t, px, py):
        """static-scene world point seen at full-resolution pixel (px, py) of frame t: ([..., 3] fp64, camera depth z [...])"""
        fx, fy, cx, cy = self.intrinsics
        px = px.to(torch.float64); py = py.to(torch.float64)
        dc = torch.stack([(px - cx) / fx, (py - cy) / fy, torch.ones_like(px)], -1)      # z = 1 => ray parameter == camera depth
        R = self.R_cw_t[t].to(px.device)
        dw = dc @ R                                                                     # R_wc dc = R_cw^T dc
        o = self.pos_t[t].to(px.device).expand_as(dw)
        z = self._cast(o, dw)
        return o + z[..., None] * dw, z

    def visible(self, frames_j, Xw, tol=0.03):
        """is world point Xw [E, 3] the first static hit along its ray from the camera centre of frame frames_j [E]?"""
        o = self.pos_t.to(Xw.device)[frames_j]
        d = Xw - o
        return (self._cast(o, d) - 1.0).abs() < tol

    def render(self, t):
        """-> image uint8 [H, W, 3] (BGR), depth fp32 [H, W], mask bool [H, W] (True = static / valid)"""
        Xw, z = self.raycast_world(t, self._px, self._py)
        r = value_noise3(Xw, self.seed * 7 + 1)
        g = value_noise3(Xw * 1.07 + 3.1, self.seed * 7 + 2)
        b = value_noise3(Xw * 0.93 - 1.7, self.seed * 7 + 3)
        shade = (0.55 + 0.45 * torch.clamp(2.5 / z.to(torch.float32), 0.0, 1.0))
        img = torch.stack([b, g, r], -1) * shade[..., None]
        img = (img - 0.25) * 1.9
        depth = z.to(torch.float32)
        mask = torch.ones(self.ht, self.wd, dtype=torch.bool, device=self.device)
        for (x0, y0, x1, y1, dd, sd) in self.dyn_rects(t):
            inside = (self._px >= x0) & (self._px <= x1) & (self._py >= y0) & (self._py <= y1)
            local = torch.stack([(self._px - x0) * 0.02, (self._py - y0) * 0.02, torch.full_like(self._px, 0.37)], -1)
            tex = value_noise3(local, sd, octaves=3, freq=2.0)
            col = torch.stack([tex, 1.0 - tex, 0.5 * tex + 0.25], -1)
            img = torch.where(inside[..., None], col, img)
            depth = torch.where(inside, torch.full_like(depth, dd), depth)
            mask = mask & ~inside
        image = (img.clamp(0.0, 1.0) * 255.0).round().to(torch.uint8)
        med = torch.median(depth[depth > 0])
        depth = torch.minimum(depth, 10.0 * med)                                        # dpvo/stream.py:63-64
        return image, depth, mask

    def frame(self, t):
        """the tuple dpvo_demo.py:107-110 passes to DPVO.__call__ after process_tensor_inputs: (t, image [3,H,W] uint8, depth, mask,
        intrinsics fp64 [4]) on self.device"""
        image, depth, mask = self.render(t)
        return t, image.permute(2, 0, 1).contiguous(), depth, mask, torch.tensor(self.intrinsics, device=self.device)


def umeyama_sim3(src, dst):
    """least-squares Sim(3) (s, R, t) with dst ~ s R src + t (Umeyama 1991); src, dst [n, 3] fp64"""
    mu_s, mu_d = src.mean(0), dst.mean(0)
    xs, xd = src - mu_s, dst - mu_d
    cov = xd.T @ xs / len(src)
    U, D, Vt = np.linalg.svd(cov)
    S = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        S[2, 2] = -1
    R = U @ S @ Vt
    var = (xs ** 2).sum() / len(src)
    s = np.trace(np.diag(D) @ S) / var
    return s, R, mu_d - s * R @ mu_s


def ate_rmse(est_xyz, gt_xyz):
    """translation ATE-RMSE after Sim(3) alignment -- the recipe of evaluation/dpvo_eva_tum.py:174-176 (evo APE, translation
    part, align=True, correct_scale=True)"""
    est_xyz = np.asarray(est_xyz, np.float64); gt_xyz = np.asarray(gt_xyz, np.float64)
    s, R, t = umeyama_sim3(est_xyz, gt_xyz)
    err = (s * est_xyz @ R.T + t) - gt_xyz
    return float(np.sqrt((err ** 2).sum(-1).mean()))
