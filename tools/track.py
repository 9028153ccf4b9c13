#!/usr/bin/env python
"""Run the reference's unchanged tracker on the synthetic clip with one backend (one per process) -- measurement / test tool.

    python tools/track.py --backend b200    --frames 200 [--M 384] [--log] [--calls] [--out gpurun_out/track_b200.json]
    python tools/track.py --backend refcuda --frames 200 ...

--log    per-frame (n, m, E, crc(ii), crc(jj), crc(kk), initialised) after every DPVO.__call__ (one host sync per frame)
--calls  (b200 only) teacher-forced per-call comparison: every altcorr.corr, fastba.neighbors and fastba.BA call the tracker makes
         is repeated on clones of the same inputs by the reference's own CUDA build (oracle/_ref/refgpu_*.so, loaded under alias
         module names) and the results are compared; the worst deviations are reported.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from harness import tracker as H  # noqa: E402
from harness.world import World  # noqa: E402


def quat_angle(qa, qb):
    d = (qa * qb).sum(-1).abs().clamp(max=1.0)
    v = qa[..., 3:] * qb[..., :3] - qb[..., 3:] * qa[..., :3] - torch.linalg.cross(qa[..., :3], qb[..., :3])
    return 2 * torch.atan2(v.norm(dim=-1), d)


class CallComparer:
    """wraps dpvo.altcorr.corr / dpvo.fastba.neighbors / dpvo.fastba.BA (module attributes the unchanged host code looks up at
    call time) so that each call also runs on the reference's CUDA build with identical inputs"""

    def __init__(self):
        from harness import _refgpu
        self.rc, self.rb = _refgpu.load("cuda_corr"), _refgpu.load("cuda_ba")
        self.stats = dict(corr_calls=0, corr_max_abs=0.0, corr_max_rel_to_peak=0.0, nb_calls=0, nb_mismatch=0, ba_calls=0,
                          ba_max_dt=0.0, ba_max_dq_rad=0.0, ba_max_dd_rel=0.0, ba_max_E=0, ba_calls_compared=0,
                          ref_self_max_dt=0.0, ref_self_max_dq_rad=0.0, ref_self_max_dd_rel=0.0, ba_per_call=[])

    def install(self):
        import dpvo.altcorr as A
        import dpvo.fastba as F
        ours_corr, ours_nb, ours_BA = A.corr, F.neighbors, F.BA
        st = self.stats

        def corr(fmap1, fmap2, coords, ii, jj, radius=1, dropout=1):
            out = ours_corr(fmap1, fmap2, coords, ii, jj, radius, dropout)
            ref, = self.rc.forward(fmap1, fmap2, coords, ii, jj, radius)
            ref = ref.float()
            d = (out.float() - ref).abs().max().item() if ref.numel() else 0.0
            st["corr_calls"] += 1
            st["corr_max_abs"] = max(st["corr_max_abs"], d)
            st["corr_max_rel_to_peak"] = max(st["corr_max_rel_to_peak"], d / max(1.0, ref.abs().max().item() if ref.numel() else 1.0))
            return out

        def neighbors(a, b):
            ix, jx = ours_nb(a, b)
            rix, rjx = self.rb.neighbors(a, b)
            st["nb_calls"] += 1
            st["nb_mismatch"] += int(not (torch.equal(ix, rix) and torch.equal(jx, rjx)))
            return ix, jx

        def BA(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, iterations=2):
            p_ref, d_ref = poses.data.clone(), patches.clone()
            p_ref2, d_ref2 = poses.data.clone(), patches.clone()
            self.rb.forward(p_ref, d_ref, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, iterations)
            self.rb.forward(p_ref2, d_ref2, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, iterations)   # its own atomics spread
            r = ours_BA(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, iterations)
            st["ba_calls"] += 1
            P, Pr = poses.data.view(-1, 7)[t0:t1], p_ref.view(-1, 7)[t0:t1]
            if P.numel() and torch.isfinite(Pr).all():
                st["ba_calls_compared"] += 1
                st["ba_max_dt"] = max(st["ba_max_dt"], (P[:, :3] - Pr[:, :3]).abs().max().item())
                st["ba_max_dq_rad"] = max(st["ba_max_dq_rad"], quat_angle(P[:, 3:], Pr[:, 3:]).max().item())
                ks = torch.unique(kk)
                d, dr = patches.view(-1, 3, 3, 3)[ks, 2, 1, 1], d_ref.view(-1, 3, 3, 3)[ks, 2, 1, 1]
                st["ba_max_dd_rel"] = max(st["ba_max_dd_rel"], ((d - dr).abs() / dr.abs().clamp(min=1e-2)).max().item())
                st["ba_max_E"] = max(st["ba_max_E"], int(ii.numel()))
                Pr2 = p_ref2.view(-1, 7)[t0:t1]
                dr2 = d_ref2.view(-1, 3, 3, 3)[ks, 2, 1, 1]
                rec = dict(call=st["ba_calls"], t0=int(t0), t1=int(t1), E=int(ii.numel()),
                           dt=(P[:, :3] - Pr[:, :3]).abs().max().item(), dq=quat_angle(P[:, 3:], Pr[:, 3:]).max().item(),
                           dd=((d - dr).abs() / dr.abs().clamp(min=1e-2)).max().item(),
                           dd_med=((d - dr).abs() / dr.abs().clamp(min=1e-2)).median().item(),
                           ref_dt=(Pr2[:, :3] - Pr[:, :3]).abs().max().item(), ref_dq=quat_angle(Pr2[:, 3:], Pr[:, 3:]).max().item(),
                           ref_dd=((dr2 - dr).abs() / dr.abs().clamp(min=1e-2)).max().item(),
                           tmag=Pr[:, :3].abs().max().item())
                st["ref_self_max_dt"] = max(st["ref_self_max_dt"], rec["ref_dt"])
                st["ref_self_max_dq_rad"] = max(st["ref_self_max_dq_rad"], rec["ref_dq"])
                st["ref_self_max_dd_rel"] = max(st["ref_self_max_dd_rel"], rec["ref_dd"])
                st["ba_per_call"].append(rec)
            return r

        A.corr, F.neighbors, F.BA = corr, neighbors, BA


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="b200", choices=["b200", "refcuda"])
    ap.add_argument("--frames", type=int, default=200)
    ap.add_argument("--M", type=int, default=384)
    ap.add_argument("--yaml", default="default.yaml")
    ap.add_argument("--seed", type=int, default=2)
    ap.add_argument("--ht", type=int, default=384)
    ap.add_argument("--wd", type=int, default=512)
    ap.add_argument("--log", action="store_true")
    ap.add_argument("--calls", action="store_true")
    ap.add_argument("--host-frames", action="store_true")
    ap.add_argument("--no-depth", action="store_true")
    ap.add_argument("--no-mask", action="store_true")
    ap.add_argument("--global-ba", action="store_true")
    ap.add_argument("--fused-softagg", action="store_true", help="b200 only: dpvo.net's SoftAgg modules -> this repository's blocks.SoftAgg")
    ap.add_argument("--all-switches", action="store_true", help="b200 only: + DPVO.corr as one corr_stacked call, dpvo.dpvo.pops -> the fused projective_ops")
    ap.add_argument("--buffer", type=int, default=0)
    ap.add_argument("--out", default="")
    ap.add_argument("--save-poses", default="")
    ap.add_argument("--breakdown", action="store_true", help="per-phase ms/frame with synchronisations (attribution run)")
    ap.add_argument("--profile-frames", type=int, default=30)
    ap.add_argument("--profile", default="", help="write a per-kernel CUDA time table of the steady-state frames to this file")
    args = ap.parse_args()

    H.load_tracker(args.backend)
    over = dict(PATCHES_PER_FRAME=args.M)
    if args.buffer:
        over["BUFFER_SIZE"] = args.buffer
    if args.global_ba:
        over["ENABLE_GLOBAL_BA"] = True
    cfg = H.make_cfg(args.yaml, **over)
    world = World(seed=args.seed, ht=args.ht, wd=args.wd, n_frames=args.frames, device="cuda")
    cmp_ = None
    if args.calls:
        assert args.backend == "b200"
        cmp_ = CallComparer()
        cmp_.install()
    frames = [world.frame(t) for t in range(args.frames)]
    torch.cuda.synchronize()
    prof_state = {}

    def hooks(slam):   # profile only a window of steady-state frames: wrap __call__ through the class the instance dispatches on
        if not args.profile:
            return
        from torch.profiler import ProfilerActivity, profile
        lo, hi = args.frames - args.profile_frames - 5, args.frames - 5
        orig = type(slam).__call__

        def call(self, t, *a, **k):
            if t == lo:
                prof_state["p"] = profile(activities=[ProfilerActivity.CUDA])
                prof_state["p"].__enter__()
            r_ = orig(self, t, *a, **k)
            if t == hi - 1:
                torch.cuda.synchronize()
                prof_state["p"].__exit__(None, None, None)
            return r_
        type(slam).__call__ = call

    r = H.run_clip(world, backend=args.backend, n_frames=args.frames, cfg=cfg, log=args.log, frames_on_host=args.host_frames,
                   use_depth=not args.no_depth, use_mask=not args.no_mask, pre_rendered=frames, hooks=hooks, breakdown=args.breakdown,
                   fused_softagg="all" if args.all_switches else args.fused_softagg)
    if args.profile:
        ev = prof_state["p"].key_averages()
        rows = sorted(((e.key, e.count, e.device_time_total / 1e3) for e in ev if e.device_time_total > 0), key=lambda x: -x[2])
        tot = sum(x[2] for x in rows)
        os.makedirs(os.path.dirname(os.path.abspath(args.profile)), exist_ok=True)
        with open(args.profile, "w") as f:
            f.write(f"# {args.backend}, M={args.M}: {args.profile_frames} steady-state frames profiled (CUDA activities only); kernel time "
                    f"{tot:.1f} ms = {tot / args.profile_frames:.3f} ms/frame; whole run {r['fps']:.2f} frames/s with the profiler window on\n")
            for k, c, t in rows[:60]:
                f.write(f"{t / args.profile_frames:9.4f} ms/frame  {c / args.profile_frames:8.1f} x/frame  {k[:150]}\n")
    res = dict(backend=args.backend, frames=r["frames"], seconds=r["seconds"], fps=r["fps"], device_ms=r["device_ms"], ate=r["ate"],
               keyframes=r["keyframes"], edges=r["edges"], M=args.M, lietorch=H._loaded["lietorch_file"],
               native=[m.__file__ for m in H._loaded["native"][:2]], frame_log=r["frame_log"],
               ms_breakdown=r["ms_breakdown"], calls=(cmp_.stats if cmp_ else None), n_poses=int(len(r["poses"])), finite=bool(np.isfinite(r["poses"]).all()))
    if args.save_poses:
        np.savez(args.save_poses, poses=r["poses"], tstamps=r["tstamps"], gt=world.gt_cam_positions())
    s = json.dumps(res)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            f.write(s + "\n")
    res.pop("frame_log")
    if res["calls"]:
        res["calls"] = {k: v for k, v in res["calls"].items() if k != "ba_per_call"}
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
