"""Drives the reference's UNCHANGED tracker (baseline/_ref/dpvo/dpvo.py, installed by baseline/install_ref.py) on the synthetic
clip of harness/world.py -- TEST / MEASUREMENT INFRASTRUCTURE, not product code.

Two arms, one per process (the native modules are imported by name, dpvo/altcorr/correlation.py:2, dpvo/fastba/ba.py:2,
dpvo/lietorch/group_ops.py:1, so a process can hold only one set):

    backend "b200"     cuda_corr / cuda_ba / lietorch_backends of this repository (wild-video-3d-reconstruction_b200/*.so)
    backend "refcuda"  the reference's own .cu/.cpp compiled by nvcc for sm_100 (oracle/_ref/refgpu_*.so, harness/backends_refcuda)

Absent third-party packages (yacs, torch_scatter, rerun, matplotlib) come from harness/shims for BOTH arms.

The network: there are no pretrained weights on this box (checkpoints/dpvo.pth is absent, no network access), so the harness
passes DPVO a module instance (dpvo/dpvo.py:228-229) -- `TeacherNet`: a seeded random-init VONet that runs its full patchify and
update operator (so every frame pays the real network cost and consumes the correlation volume), whose flow head output is then
replaced by the ground-truth flow of the synthetic world plus seeded pixel noise, and its weight head by a visibility-gated
constant.  This keeps the closed loop meaningful (ATE can be measured) and identical between the arms.
"""
import importlib
import os
import sys
import time
import zlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG = "wild-video-3d-reconstruction_b200"
REF_INSTALL = os.path.join(ROOT, "baseline", "_ref")
_loaded = {}


def _have(mod):
    try:
        importlib.import_module(mod)
        return True
    except Exception:
        return False


def load_tracker(backend="b200"):
    """-> dict(dpvo=<module dpvo.dpvo>, net=<module dpvo.net>, cfg=<CfgNode defaults>, backend=...).  One backend per process."""
    if _loaded:
        assert _loaded["backend"] == backend, f"this process already runs the tracker on {_loaded['backend']}"
        return _loaded
    assert os.path.isdir(os.path.join(REF_INSTALL, "dpvo")), "reference host code missing: run `python baseline/install_ref.py`"
    shim_dir = os.path.join(HERE, "shims")
    for name in ("yacs", "torch_scatter", "rerun", "matplotlib"):
        if name not in sys.modules and not _have(name):
            if shim_dir not in sys.path:
                sys.path.append(shim_dir)          # at the END: a real installation always wins
            importlib.invalidate_caches()
            importlib.import_module(name)
    if backend == "b200":
        pkg = importlib.import_module(PKG)
        pkg.install_as_dpvo_backends()
        native_dir = os.path.dirname(pkg.__file__)
        if os.environ.get("WVB200_LIETORCH", "") == "ref":   # bring-up switch: the reference's lietorch build under this repo's corr/BA
            sys.path.insert(0, os.path.join(HERE, "backends_reflie"))
    elif backend == "refcuda":
        native_dir = os.path.join(HERE, "backends_refcuda")
        sys.path.insert(0, native_dir)
    else:
        raise ValueError(backend)
    if REF_INSTALL not in sys.path:
        sys.path.insert(0, REF_INSTALL)
    import cuda_ba
    import cuda_corr
    import lietorch_backends
    for m in (cuda_corr, cuda_ba):
        assert os.path.dirname(os.path.abspath(m.__file__)) == os.path.abspath(native_dir), (m.__file__, native_dir)
    dpvo_mod = importlib.import_module("dpvo.dpvo")
    assert os.path.abspath(dpvo_mod.__file__).startswith(REF_INSTALL), dpvo_mod.__file__
    _loaded.update(backend=backend, dpvo=dpvo_mod, net=importlib.import_module("dpvo.net"),
                   cfg=importlib.import_module("dpvo.config").cfg, lietorch_file=getattr(lietorch_backends, "__file__", "?"),
                   native=(cuda_corr, cuda_ba, lietorch_backends))
    return _loaded


class host_switches:
    """INTEGRATION.md section 1, the OPTIONAL edits of the reference's Python, applied as patches for one run (b200 backend only) and
    undone afterwards: `corr` -- DPVO.corr (dpvo/dpvo.py:325-333) as one altcorr.corr_stacked call; `pops` -- dpvo.dpvo's `pops`
    module (transform / flow_mag / point_cloud, dpvo/projective_ops.py) -> this repository's fused projective_ops.  (The SoftAgg
    switch lives in TeacherNet: it changes how the network is built.)"""

    def __init__(self, which):
        self.which = set(which or ())
        self.saved = []

    def __enter__(self):
        if not self.which:
            return self
        assert _loaded.get("backend") == "b200", "the optional switches replace reference code by this repository's"
        dm = _loaded["dpvo"]
        if "corr" in self.which:
            altcorr = importlib.import_module(PKG + ".altcorr")

            def corr(slam, coords, indicies=None):
                ii, jj = indicies if indicies is not None else (slam.pg.kk, slam.pg.jj)
                return altcorr.corr_stacked(slam.gmap, slam.pyramid, coords, ii % (slam.M * slam.pmem), jj % slam.pmem, 3)
            self.saved.append((dm.DPVO, "corr", dm.DPVO.corr))
            dm.DPVO.corr = corr
        if "pops" in self.which:
            self.saved.append((dm, "pops", dm.pops))
            dm.pops = importlib.import_module(PKG + ".projective_ops")
        return self

    def __exit__(self, *exc):
        for obj, name, old in reversed(self.saved):
            setattr(obj, name, old)
        self.saved = []
        return False


def make_cfg(yaml_name="default.yaml", **overrides):
    cfg = load_tracker(_loaded.get("backend", "b200"))["cfg"].clone()
    cfg.merge_from_file(os.path.join(REF_INSTALL, "dpvo_configs", yaml_name))
    for k, v in overrides.items():
        cfg[k] = v
    return cfg


class TeacherNet(torch.nn.Module):
    """see the module docstring; `world` supplies the ground truth"""

    def __init__(self, world, seed=1234, noise_px=0.25, weight=0.6, patches_per_frame=384, fused_softagg=False):
        super().__init__()
        net_mod = _loaded["net"]
        g = torch.random.get_rng_state()
        torch.manual_seed(seed)
        if fused_softagg:   # True or "all"
            # INTEGRATION.md section 1, optional switch: the update operator's two SoftAgg modules (dpvo/net.py:17, :44-45) become this
            # repository's blocks.SoftAgg (same parameter names and creation order, so the same seed gives the same weights): group
            # plan + one fused scatter-softmax / scatter-sum kernel instead of torch.unique + torch_scatter (shimmed in this image)
            import importlib
            ours = importlib.import_module("wild-video-3d-reconstruction_b200.blocks").SoftAgg
            ref_cls, net_mod.SoftAgg = net_mod.SoftAgg, ours
            try:
                self.base = net_mod.VONet()
            finally:
                net_mod.SoftAgg = ref_cls
        else:
            self.base = net_mod.VONet()
        torch.random.set_rng_state(g)
        self.P, self.DIM, self.RES = self.base.P, self.base.DIM, self.base.RES
        self.world = world
        self.noise_px, self.weight = float(noise_px), float(weight)
        self.slam = None
        self.frame = 0
        self.M = patches_per_frame
        self.gen = None
        self.seed = seed

    def bind(self, slam):
        self.slam = slam
        dev = slam.pg.poses_.device
        self.Xw = torch.zeros(self.world.n_frames, slam.M, 3, dtype=torch.float64, device=dev)
        self.R = self.world.R_cw_t.to(dev)
        self.t = self.world.t_cw_t.to(dev)
        self.gen = torch.Generator(device=dev)
        self.gen.manual_seed(self.seed)
        rects = np.array([[r[:4] for r in self.world.dyn_rects(t)] for t in range(self.world.n_frames)], np.float64)
        self.rects = torch.from_numpy(rects.reshape(self.world.n_frames, -1, 4) / self.RES).to(dev)   # [frames, n_dyn, (x0, y0, x1, y1)] at 1/4 res

    # -- the two entry points DPVO calls (dpvo/dpvo.py:783, :379, :720) ---------------------------------------------------------
    def patchify(self, image, **kw):
        out = self.base.patchify(image, **kw)
        patches = out[3]                                           # [1, M, 3, P, P]: x, y at 1/4 resolution, inverse depth
        c = self.P // 2
        xq, yq = patches[0, :, 0, c, c], patches[0, :, 1, c, c]
        Xw, _ = self.world.raycast_world(self.frame, self.RES * xq, self.RES * yq)
        self.Xw[self.frame] = Xw
        return out

    def update(self, net, ctx, corr, flow, ii, jj, kk):
        net, (d, w, _) = self.base.update(net, ctx, corr, flow, ii, jj, kk)
        slam = self.slam
        ts = torch.from_numpy(slam.pg.tstamps_).to(ii.device)      # keyframe index -> frame counter (8 * N bytes, no sync needed)
        fi, fj = ts[ii], ts[jj]
        Xw = self.Xw[fi, kk % slam.M]
        Xc = torch.einsum("eij,ej->ei", self.R[fj], Xw) + self.t[fj]
        fx, fy, cx, cy = (self.world.intrinsics / self.RES).tolist()
        Z = Xc[:, 2]
        front = Z > 0.2
        Zs = torch.where(front, Z, torch.ones_like(Z))
        gt = torch.stack([fx * Xc[:, 0] / Zs + cx, fy * Xc[:, 1] / Zs + cy], -1).to(torch.float32)
        vis = front & self.world.visible(fj, Xw)
        if self.rects.shape[1]:                                     # hidden behind a dynamic rectangle in the target frame
            r = self.rects[fj]
            gx, gy = gt[:, 0, None], gt[:, 1, None]
            vis = vis & ~((gx >= r[..., 0]) & (gx <= r[..., 2]) & (gy >= r[..., 1]) & (gy <= r[..., 3])).any(-1)
        c = self.P // 2
        coords = slam.reproject(indicies=(ii, jj, kk))[0, :, :, c, c]           # [E, 2] current estimate of the projection
        noise = self.noise_px * torch.randn(gt.shape, generator=self.gen, device=gt.device, dtype=torch.float32)
        # The update operator above ran in full (eager execution: nothing is elided) and its hidden state `net` is kept; its two
        # heads d, w are NOT mixed into the result: SoftAgg's scatter_add_ (harness/shims/torch_scatter) sums with atomics, and even
        # a 1e-7 px trace of that run-to-run noise is enough to flip thresholded decisions (keyframe test, the depth filter of the
        # point export), which would make "same clip, different sharding" comparisons meaningless.
        delta = torch.where(vis[:, None], gt + noise - coords, torch.zeros_like(gt))[None]
        wgt = (self.weight * vis.to(torch.float32))[None, :, None].expand(1, -1, 2).contiguous()
        return net, (delta, wgt, None)


class PhaseTimer:
    """per-phase wall time of the tracker with a device synchronisation on both sides of every phase (attribution, not throughput:
    the synchronisations remove the CPU/GPU overlap, so the phases sum to more than an un-instrumented frame).  Wraps attributes
    the unchanged host code looks up at call time; nothing on disk changes."""
    PHASES = ("patchify_net", "reproject", "corr", "update_operator", "teacher_gt", "ba", "point_cloud", "keyframe", "motion_probe")

    def __init__(self):
        self.t = {k: 0.0 for k in self.PHASES}
        self.n = {k: 0 for k in self.PHASES}
        self.stack = []

    def wrap(self, name, fn):
        def timed(*a, **k):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            self.stack.append(0.0)
            try:
                return fn(*a, **k)
            finally:
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                inner = self.stack.pop()
                self.t[name] += dt - inner          # exclusive time: nested phases are charged to themselves
                self.n[name] += 1
                if self.stack:
                    self.stack[-1] += dt
        return timed

    def install(self, slam, net):
        import dpvo.fastba as F
        import dpvo.projective_ops as P
        net.base.patchify.forward = self.wrap("patchify_net", net.base.patchify.forward)
        net.base.update.forward = self.wrap("update_operator", net.base.update.forward)
        net.update = self.wrap("teacher_gt", net.update)
        slam.reproject = self.wrap("reproject", slam.reproject)
        slam.corr = self.wrap("corr", slam.corr)
        slam.keyframe = self.wrap("keyframe", slam.keyframe)
        slam.motion_probe = self.wrap("motion_probe", slam.motion_probe)
        self._F, self._P, self._old = F, P, (F.BA, P.point_cloud)
        F.BA = self.wrap("ba", F.BA)
        P.point_cloud = self.wrap("point_cloud", P.point_cloud)

    def uninstall(self):
        self._F.BA, self._P.point_cloud = self._old

    def report(self, frames, total_s):
        out = {k: round(1e3 * v / max(frames, 1), 4) for k, v in self.t.items()}
        out["other_host_code"] = round(1e3 * (total_s - sum(self.t.values())) / max(frames, 1), 4)
        out["total"] = round(1e3 * total_s / max(frames, 1), 4)
        out["calls_per_frame"] = {k: round(v / max(frames, 1), 3) for k, v in self.n.items()}
        return out


def crc(t):
    return zlib.crc32(t.detach().cpu().numpy().tobytes())


def run_clip(world, backend="b200", n_frames=None, cfg=None, use_depth=True, use_mask=True, log=False, frames_on_host=False,
             seed=1234, noise_px=0.25, final_updates=12, hooks=None, pre_rendered=None, breakdown=False, fused_softagg=False):
    """the loop of dpvo_demo.py:62-141 on the synthetic clip.  Returns a dict with the trajectory (terminate()), per-frame log
    (log=True: n, m, E and CRCs of ii/jj/kk after every frame -- one D2H sync per frame, off for timing), timings, ATE."""
    T = load_tracker(backend)
    _sw = host_switches(("corr", "pops") if fused_softagg == "all" else ())
    _sw.__enter__()
    DPVO = T["dpvo"].DPVO
    cfg = cfg if cfg is not None else make_cfg()
    n_frames = n_frames or world.n_frames
    dev = torch.device("cuda")
    torch.manual_seed(seed); torch.cuda.manual_seed_all(seed); np.random.seed(seed)
    net = TeacherNet(world, seed=seed, noise_px=noise_px, patches_per_frame=cfg.PATCHES_PER_FRAME, fused_softagg=fused_softagg)
    frames = pre_rendered if pre_rendered is not None else [world.frame(t) for t in range(n_frames)]
    if frames_on_host:   # the e2e form: every frame starts in pinned host memory, dpvo_demo.py:55-60 uploads it
        frames = [(t, im.cpu().pin_memory(), dp.cpu().pin_memory(), mk.cpu().pin_memory(), K.cpu().pin_memory()) for (t, im, dp, mk, K) in frames]
    torch.cuda.synchronize()
    slam = None
    frame_log = []
    h2d = d2h = 0
    pose_host = torch.empty(7, dtype=torch.float32).pin_memory()
    t_start = time.perf_counter()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    with torch.no_grad():
        for (t, image, depth, mask, K) in frames[:n_frames]:
            if frames_on_host:
                h2d += image.numel() + depth.numel() * 4 + mask.numel() + K.numel() * 8
                image, depth, mask, K = (x.to(dev, non_blocking=True) for x in (image, depth, mask, K))
            if slam is None:
                slam = DPVO(cfg, net, ht=image.shape[1], wd=image.shape[2], viz=False)
                net.bind(slam)
                if hooks:
                    hooks(slam)
                if breakdown:
                    phases = PhaseTimer()
                    phases.install(slam, net)
            net.frame = t
            slam(t, image, depth if use_depth else None, mask if use_mask else None, K)
            if frames_on_host:   # the per-frame result a caller reads: the newest pose
                pose_host.copy_(slam.pg.poses_[max(slam.n - 1, 0)], non_blocking=False)
                d2h += 28
            if log:
                pg = slam.pg
                frame_log.append((int(slam.n), int(slam.m), int(pg.ii.numel()), crc(pg.ii), crc(pg.jj), crc(pg.kk), bool(slam.is_initialized)))
        for _ in range(final_updates):
            slam.update()
        ev1.record()
        torch.cuda.synchronize()
        t_track = time.perf_counter() - t_start
        n_kf, E_final = int(slam.n), int(slam.pg.ii.numel())
        ms_breakdown = None
        if breakdown:
            phases.uninstall()
            ms_breakdown = phases.report(n_frames, t_track)
        poses, tstamps = slam.terminate()
    est = np.asarray(poses)[:, :3]
    gt = world.gt_cam_positions()[:len(est)]
    _sw.__exit__()
    return dict(poses=np.asarray(poses), tstamps=np.asarray(tstamps), frame_log=frame_log, seconds=t_track, frames=n_frames,
                ms_breakdown=ms_breakdown, fps=n_frames / t_track, device_ms=ev0.elapsed_time(ev1), ate=ate_rmse_safe(est, gt), keyframes=n_kf, edges=E_final,
                h2d_bytes_per_frame=h2d / max(n_frames, 1), d2h_bytes_per_frame=d2h / max(n_frames, 1), slam=slam, net=net)


def ate_rmse_safe(est, gt):
    from .world import ate_rmse
    if len(est) < 3 or not np.isfinite(est).all():
        return float("nan")
    return ate_rmse(est, gt)


def run_clip_windows(world, backend, warmup, steps, breakdown_frames=0, cfg=None, frames=None, frames_on_host=False, seed=1234,
                     noise_px=0.25, barrier=None, sampler=None, net=None, fused_softagg=False):
    """bench form of run_clip: ONE tracker instance fed `warmup` untimed frames (they cover the tracker's initialisation: 10 frames
    + 12 updates, dpvo/dpvo.py:858-863), then exactly `steps` timed frames bracketed by barrier() + torch.cuda.synchronize() and a
    pair of CUDA events, then `breakdown_frames` more frames under the PhaseTimer.  frames_on_host: every frame starts in pinned host
    memory and is uploaded inside the timed region (dpvo_demo.py:55-60), and the newest pose is read back after every frame.
    -> dict(timed_ms, wall_s, h2d/d2h bytes per frame, ms_breakdown, ate, keyframes, edges, launches_per_frame)"""
    T = load_tracker(backend)
    _sw = host_switches(("corr", "pops") if fused_softagg == "all" else ())
    _sw.__enter__()
    DPVO = T["dpvo"].DPVO
    cfg = cfg if cfg is not None else make_cfg()
    dev = torch.device("cuda")
    total = warmup + steps + breakdown_frames
    assert total <= world.n_frames
    torch.manual_seed(seed); torch.cuda.manual_seed_all(seed); np.random.seed(seed)
    if net is None:
        net = TeacherNet(world, seed=seed, noise_px=noise_px, patches_per_frame=cfg.PATCHES_PER_FRAME, fused_softagg=fused_softagg)
    frames = frames if frames is not None else [world.frame(t) for t in range(total)]
    if frames_on_host:
        frames = [(t, im.cpu().pin_memory(), dp.cpu().pin_memory(), mk.cpu().pin_memory(), K.cpu().pin_memory()) for (t, im, dp, mk, K) in frames[:total]]
    sync = barrier if barrier is not None else torch.cuda.synchronize
    pose_host = torch.empty(7, dtype=torch.float32).pin_memory()
    slam, phases = None, None
    h2d = d2h = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall = 0.0
    with torch.no_grad():
        for (t, image, depth, mask, K) in frames[:total]:
            if t == warmup:
                sync()
                if sampler is not None:
                    sampler.start()
                wall = time.perf_counter()
                ev0.record()
            if t == warmup + steps:
                ev1.record()
                sync()
                wall = time.perf_counter() - wall
                if sampler is not None:
                    sampler.stop()
                if breakdown_frames:
                    phases = PhaseTimer()
                    phases.install(slam, net)
                    t_bd = time.perf_counter()
            if frames_on_host:
                if warmup <= t < warmup + steps:
                    h2d += image.numel() + depth.numel() * 4 + mask.numel() + K.numel() * 8
                image, depth, mask, K = (x.to(dev, non_blocking=True) for x in (image, depth, mask, K))
            if slam is None:
                slam = DPVO(cfg, net, ht=image.shape[1], wd=image.shape[2], viz=False)
                net.bind(slam)
            net.frame = t
            slam(t, image, depth, mask, K)
            if frames_on_host:
                pose_host.copy_(slam.pg.poses_[max(slam.n - 1, 0)], non_blocking=False)   # blocks until the frame's BA is done
                if warmup <= t < warmup + steps:
                    d2h += 28
        if breakdown_frames == 0:
            ev1.record()
            sync()
            wall = time.perf_counter() - wall
            if sampler is not None:
                sampler.stop()
        torch.cuda.synchronize()
        ms_breakdown = None
        if phases is not None:
            phases.uninstall()
            ms_breakdown = phases.report(breakdown_frames, time.perf_counter() - t_bd)
        n_kf, E_final = int(slam.n), int(slam.pg.ii.numel())
        for _ in range(12):
            slam.update()
        poses, _ = slam.terminate()
    est = np.asarray(poses)[:, :3]
    _sw.__exit__()
    return dict(timed_ms=ev0.elapsed_time(ev1), wall_s=wall, h2d_bytes_per_frame=h2d / max(steps, 1), d2h_bytes_per_frame=d2h / max(steps, 1),
                ms_breakdown=ms_breakdown, ate=ate_rmse_safe(est, world.gt_cam_positions()[:len(est)]), keyframes=n_kf, edges=E_final,
                frames_total=total, net=net)
