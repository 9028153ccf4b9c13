#!/usr/bin/env python
"""BASELINE.json configs[4]: a batch of independent synthetic clips sharded over the GPUs of one box, one process per GPU, the
clips of a rank tracked sequentially, `--export_colmap` text output per clip.  The reference's own batch runner is sequential on
one GPU (test_scripts/run_multi_scenes.py:26-63: one `dpvo_demo.py --export_colmap` subprocess per scene); clips share nothing,
so there is no data-path collective -- the only communication is the rendezvous and the max-over-ranks of the wall time.

    python tools/run_clips.py --clips 8 --frames 120 --out gpurun_out/clips_1gpu                               # one GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \\
        tools/run_clips.py --clips 64 --frames 4500 --out gpurun_out/clips_8gpu                                # 8 GPUs, 8 clips each

Per clip c (world seed 100 + c) the loop is dpvo_demo.py:62-141 + :186-208: every frame through the UNCHANGED
DPVO.__call__ on this repository's native modules, 12 final updates, get_pts_clr_intri(), terminate(), then
save_output_for_COLMAP (this repository's byte-identical mirror of dpvo/plot_utils.py:58-115) into <out>/clip_<c>/colmap/.
Rank 0 prints ONE JSON line: clips, frames, wall seconds (max over ranks), aggregate frames/s, per-clip ATE.

`--compare A B` checks two output trees against each other the way SURVEY 8d config 5 states it: images.txt ID / camera / name
columns byte-identical, pose and point columns within 1e-4.
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "wild-video-3d-reconstruction_b200"


def track_clip(c, args, H, World, colmap, torch):
    world = World(seed=100 + c, n_frames=args.frames, device="cuda", ht=args.ht, wd=args.wd)
    cfg = H.make_cfg(args.yaml, PATCHES_PER_FRAME=args.M, BUFFER_SIZE=args.buffer)
    torch.manual_seed(1234 + c)                        # patch sampling is the tracker's only randomness: fixed per clip, not per rank
    r = H.run_clip(world, backend="b200", n_frames=args.frames, cfg=cfg, seed=1234 + c)
    slam = r["slam"]
    poses, tstamps = r["poses"], r["tstamps"]
    points, colors, (intr, Hh, Ww) = slam.get_pts_clr_intri(inlier=False)      # dpvo_demo.py:137 (inlier=True only adds a PNG plot)
    out = os.path.join(args.out, f"clip_{c:03d}")
    os.makedirs(out, exist_ok=True)
    traj = colmap.Trajectory(positions_xyz=poses[:, :3], orientations_quat_wxyz=poses[:, [6, 3, 4, 5]], timestamps=tstamps)   # dpvo_demo.py:187-191
    colmap.save_output_for_COLMAP(os.path.join(out, "colmap"), tstamps, traj, points, colors, False, *intr, Hh, Ww)   # dpvo_demo.py:205
    del slam, r["slam"], r["net"]
    torch.cuda.empty_cache()
    return dict(clip=c, frames=args.frames, seconds=r["seconds"], ate=r["ate"], keyframes=r["keyframes"], points=int(len(points)))


def read_images_txt(path):
    rows = []
    with open(path) as f:
        for line in f:
            p = line.split()
            if len(p) == 10:
                rows.append((p[0], p[8], p[9], [float(v) for v in p[1:8]]))
    return rows


def compare_trees(a, b, tol=1e-4, subset=False):
    """-> dict(clips, max_pose_diff, max_point_diff); raises AssertionError on the first violation.  subset: compare the clips
    of tree b only (b = a re-run of some clips of a under another sharding)"""
    clips = sorted(d for d in os.listdir(b if subset else a) if d.startswith("clip_"))
    if subset:
        assert clips and all(os.path.isdir(os.path.join(a, c)) for c in clips), "clips of the second tree missing from the first"
    else:
        assert clips and clips == sorted(d for d in os.listdir(b) if d.startswith("clip_")), "different clip sets"
    worst_pose = worst_pt = 0.0
    for c in clips:
        ia, ib = read_images_txt(os.path.join(a, c, "colmap", "images.txt")), read_images_txt(os.path.join(b, c, "colmap", "images.txt"))
        assert len(ia) == len(ib) and len(ia) > 0, (c, len(ia), len(ib))
        for ra, rb in zip(ia, ib):
            assert ra[:3] == rb[:3], (c, ra[:3], rb[:3])                      # IMAGE_ID, CAMERA_ID, NAME: byte-identical
            worst_pose = max(worst_pose, float(np.abs(np.array(ra[3]) - np.array(rb[3])).max()))
        assert open(os.path.join(a, c, "colmap", "cameras.txt")).read() == open(os.path.join(b, c, "colmap", "cameras.txt")).read()
        pa = np.loadtxt(os.path.join(a, c, "colmap", "points3D.txt"), ndmin=2)
        pb = np.loadtxt(os.path.join(b, c, "colmap", "points3D.txt"), ndmin=2)
        assert pa.shape == pb.shape, (c, pa.shape, pb.shape)
        if pa.size:
            assert np.array_equal(pa[:, 0], pb[:, 0]) and np.array_equal(pa[:, 4:], pb[:, 4:]), c   # ids, colours, trailing columns
            worst_pt = max(worst_pt, float(np.abs(pa[:, 1:4] - pb[:, 1:4]).max()))
    assert worst_pose <= tol, f"pose columns differ by {worst_pose}"
    assert worst_pt <= tol, f"point columns differ by {worst_pt}"
    return dict(clips=len(clips), max_pose_diff=worst_pose, max_point_diff=worst_pt)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clips", type=int, default=8)
    ap.add_argument("--frames", type=int, default=120, help="frames per clip (configs[4]: 4500 = a 15-minute clip at 5 FPS)")
    ap.add_argument("--M", type=int, default=384)
    ap.add_argument("--yaml", default="default.yaml")
    ap.add_argument("--buffer", type=int, default=2048)
    ap.add_argument("--ht", type=int, default=384)
    ap.add_argument("--wd", type=int, default=512)
    ap.add_argument("--out", default="gpurun_out/clips")
    ap.add_argument("--compare", nargs=2, default=None)
    ap.add_argument("--subset", action="store_true", help="--compare: the second tree holds a subset of the first tree's clips")
    ap.add_argument("--only", default=None, help="comma-separated clip ids to track instead of 0 .. clips-1 (sharded the same way)")
    args = ap.parse_args()
    if args.compare:
        print(json.dumps(compare_trees(*args.compare, subset=args.subset)))
        return 0
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import torch
    dist_mod = importlib.import_module(PKG + ".dist")
    backend = os.environ.get("WV_DIST_BACKEND") or None       # "gloo" lets several ranks share one GPU (single-GPU test boxes)
    rank, world, local = dist_mod.init(backend)
    torch.cuda.set_device(local % torch.cuda.device_count())
    from harness import tracker as H
    from harness.world import World
    H.load_tracker("b200")
    colmap = importlib.import_module(PKG + ".colmap_export")
    ids = [int(c) for c in args.only.split(",")] if args.only else list(range(args.clips))
    mine = [ids[i] for i in dist_mod.clip_shard(len(ids), rank, world)]
    args.clips = len(ids)
    os.makedirs(args.out, exist_ok=True)
    dist_mod.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    done = [track_clip(c, args, H, World, colmap, torch) for c in mine]
    torch.cuda.synchronize()
    t_mine = time.perf_counter() - t0
    t_max = dist_mod.max_over_ranks(t_mine, device="cpu" if backend == "gloo" else "cuda")
    with open(os.path.join(args.out, f"rank_{rank:02d}.json"), "w") as f:
        json.dump(done, f)
    dist_mod.barrier()
    if rank == 0:
        allc = []
        for r in range(world):
            with open(os.path.join(args.out, f"rank_{r:02d}.json")) as f:
                allc += json.load(f)
        line = {"metric": "clip_batch_frames_per_s", "value": args.clips * args.frames / t_max, "unit": "frames/s", "n_gpus": world,
                "clips": args.clips, "frames_per_clip": args.frames, "seconds": t_max, "clips_per_rank": [len(dist_mod.clip_shard(args.clips, r, world)) for r in range(world)],
                "config": {"workload": f"configs[4]: {args.clips} synthetic clips x {args.frames} frames (seeds 100..), {args.yaml} M={args.M}, --export_colmap text output",
                           "parallelism": f"clip-level data parallelism over {world} process(es), sequential per rank, no collective"},
                "ate_rmse_m": {"median": float(np.median([c["ate"] for c in allc])), "max": float(np.max([c["ate"] for c in allc]))},
                "per_clip": sorted(allc, key=lambda c: c["clip"])}
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    dist_mod.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
