"""Global BA (BASELINE config 4: long buffer, all poses but the first free), single GPU or patch-sharded under torchrun.

    python tools/global_ba.py --frames 1024                      # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/global_ba.py --frames 1024 --check                 # 2 GPUs, NCCL allreduce of [S | y] and dZ

--check: (a) all ranks end with bitwise identical poses / patches, (b) the sharded result equals the one-GPU cuda_ba.forward
result to rounding.  Prints one JSON line on rank 0 (device time, max over ranks).  The edge list is the
pattern of the reference's global_bundle_adjustment (dpvo/dpvo.py:448-471): (i, i+1) plus (i, i+10..i+19) for every 5th frame.
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=512)
    ap.add_argument("--M", type=int, default=96)
    ap.add_argument("--iters", type=int, default=2)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--edges", default="global", help="global (dpvo.py:448-471 pattern) | tracker | all")
    ap.add_argument("--loop-frac", type=float, default=0.02, help="config4: share of extra edges to random frames")
    a = ap.parse_args()
    rank, world, local = pkg.dist.init()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _, cb = pkg.native_modules()
    fb = importlib.import_module(pkg.__name__ + ".fastba")
    s = pkg.synthetic.make_ba_scene(seed=5, n=a.frames, M=a.M, edges=a.edges, pose_noise=0.005, loop_frac=a.loop_frac)
    D = {k: torch.from_numpy(np.ascontiguousarray(s[k])).to(dev) for k in
         ("poses", "patches", "intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")}
    t0, t1 = 1, a.frames
    E = D["ii"].numel()

    def run(sharded):
        poses, patches = D["poses"].clone(), D["patches"].clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        pkg.dist.barrier(); torch.cuda.synchronize()
        e0.record()
        if sharded:
            fb.BA_sharded(poses, patches, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, a.iters)
        else:
            cb.forward(poses, patches, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, a.iters)
        e1.record(); torch.cuda.synchronize()
        return pkg.dist.max_over_ranks(e0.elapsed_time(e1), device=dev), poses, patches

    times = []
    for i in range(a.reps + 2):
        ms, poses, patches = run(True)
        if i >= 2:
            times.append(ms)
    out = {"what": "global BA, %d GN iterations" % a.iters, "n_gpus": world, "frames": a.frames, "free_poses": t1 - t0, "edges": E,
           "ms": float(np.median(times)), "edges_per_s": E * a.iters / (float(np.median(times)) * 1e-3)}
    # per-piece device times of one iteration on this rank's shard
    wp, wq = D["poses"].clone(), D["patches"].clone()
    args = (wp, wq, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    cb.global_build(*args, rank, world, True); torch.cuda.synchronize()
    ev[0].record(); Sy = cb.global_build(*args, rank, world, True)
    ev[1].record(); dX = cb.global_solve(Sy, t1 - t0)
    ev[2].record(); dZ = cb.global_dz(*args, rank, world, dX)
    ev[3].record(); cb.global_retract(wp, wq, E, t0, t1, dX, dZ)
    ev[4].record(); Sy = cb.global_build(*args, rank, world, False)
    ev[5].record(); torch.cuda.synchronize()
    out["pieces_ms"] = {k: round(ev[i].elapsed_time(ev[i + 1]), 4)
                        for i, k in enumerate(("plan+build", "damp+cholesky+solve", "dz", "retract", "build (plan reused)"))}
    del Sy
    if a.check:
        _, p1, q1 = run(False)
        out["max_dpose_vs_one_gpu"] = float((poses - p1).abs().max())
        out["max_ddepth_vs_one_gpu"] = float((patches - q1).abs().max())
        if world > 1:
            import torch.distributed as dist
            ref = [torch.empty_like(poses) for _ in range(world)]
            dist.all_gather(ref, poses)
            out["ranks_identical"] = bool(all(torch.equal(r, ref[0]) for r in ref))
        err = float(np.abs(poses.cpu().numpy()[:, :3] - s["poses_gt"][:, :3]).max())
        err0 = float(np.abs(s["poses"][:, :3] - s["poses_gt"][:, :3]).max())
        out["max_pos_err_before"], out["max_pos_err_after"] = err0, err
    if rank == 0:
        print(json.dumps(out), flush=True)
    pkg.dist.finalize()


if __name__ == "__main__":
    main()
