"""Global BA (BASELINE configs[3]: long buffer, all poses but the first free), single GPU or patch-sharded under torchrun.

    python tools/global_ba.py --frames 1024                      # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/global_ba.py --frames 1024 --check                 # 2 GPUs, NCCL allreduce of [S | y] and dZ
    python tools/global_ba.py --probe                            # band solve vs dense solve on the same reduced systems

--check: (a) all ranks end with bitwise identical poses / patches, (b) the sharded result against the one-GPU cuda_ba.forward
result.  Prints one JSON line on rank 0 (device time, max over ranks).  --edges global is the pattern of the reference's
global_bundle_adjustment (dpvo/dpvo.py:448-471): (i, i+1) plus (i, i+10..i+19) for every 5th frame (banded, span 19);
--edges config4 is SURVEY 8d config 4 (+-12 frames per patch; --loop-frac > 0 adds edges to random frames, which makes the reduced
system dense and takes the dense path).  WVB200_GLOBAL_DENSE=1 forces the dense path.

bench_sharded() is the "sharded" sub-object of bench.py's JSON line.
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
KEYS = ("poses", "patches", "intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")


def to_dev(s, dev):
    return {k: torch.from_numpy(np.ascontiguousarray(s[k])).to(dev) for k in KEYS}


def cost(D, poses, patches):
    """weighted squared reprojection error of the edge list (fastba's projection: row-0 intrinsics, no depth clamp), fp64, gauge-free"""
    _, cb = pkg.native_modules()
    co = cb.reproject(poses, patches, D["intrinsics"], D["ii"], D["jj"], D["kk"])[0, :, :, 1, 1].double()
    r = D["target"].double() - co
    ok = torch.isfinite(r).all(-1) & (r.norm(dim=-1) < 128)
    return float((D["weight"].double() * r * r)[ok].sum())


def run_sharded(D, t0, t1, iters, sharded=True):
    """-> (ms max over ranks, poses, patches)"""
    _, cb = pkg.native_modules()
    fb = importlib.import_module(pkg.__name__ + ".fastba")
    poses, patches = D["poses"].clone(), D["patches"].clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pkg.dist.barrier(); torch.cuda.synchronize()
    e0.record()
    if sharded:
        fb.BA_sharded(poses, patches, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, iters)
    else:
        cb.forward(poses, patches, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, iters)
    e1.record(); torch.cuda.synchronize()
    return pkg.dist.max_over_ranks(e0.elapsed_time(e1), device=poses.device), poses, patches


def pieces(D, t0, t1, rank, world):
    """per-piece device times (ms) of ONE iteration on this rank's shard + the bytes one all-reduce of [S | y] moves"""
    _, cb = pkg.native_modules()
    wp, wq = D["poses"].clone(), D["patches"].clone()
    args = (wp, wq, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    span = cb.global_span(wp, wq, D["ii"], D["jj"], D["kk"], t0, t1, True)
    bs = cb.band_block(span)
    build = (lambda: cb.global_build_band(*args, rank, world, bs)) if bs else (lambda: cb.global_build(*args, rank, world, False))
    solve = (lambda Sy: cb.global_solve_band(Sy, t1 - t0, bs)) if bs else (lambda Sy: cb.global_solve(Sy, t1 - t0))
    Sy = build(); dX = solve(Sy); torch.cuda.synchronize()     # warm
    ev[0].record(); span = cb.global_span(wp, wq, D["ii"], D["jj"], D["kk"], t0, t1, True)
    ev[1].record(); Sy = build()
    ev[2].record(); dX = solve(Sy)
    ev[3].record(); dZ = cb.global_dz(*args, rank, world, dX)
    ev[4].record(); cb.global_retract(wp, wq, D["ii"].numel(), t0, t1, dX, dZ)
    ev[5].record(); torch.cuda.synchronize()
    out = {k: round(ev[i].elapsed_time(ev[i + 1]), 4) for i, k in enumerate(("plan+span", "build", "solve", "dz", "retract"))}
    return out, int(span), int(bs), int(Sy.numel() * 4)


def probe(dev):
    """band solve vs dense solve of the SAME reduced system (both built by ba_global_kernel from the same inputs), several sizes"""
    _, cb = pkg.native_modules()
    rows = []
    for (n, M, edges) in ((24, 16, "global"), (64, 32, "global"), (200, 24, "config4"), (600, 8, "global")):
        s = pkg.synthetic.make_ba_scene(seed=7, n=n, M=M, edges=edges, pose_noise=0.005, loop_frac=0.0)
        D = to_dev(s, dev)
        t0, t1 = 1, n
        args = (D["poses"].clone(), D["patches"].clone(), D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1)
        span = cb.global_span(args[0], args[1], D["ii"], D["jj"], D["kk"], t0, t1, True)
        bs = cb.band_block(span)
        n6 = 6 * (t1 - t0)
        Sd = cb.global_build(*args, 0, 1, False).clone()
        Sb = cb.global_build_band(*args, 0, 1, bs)
        dXd = cb.global_solve(Sd.clone(), t1 - t0).double()
        dXb = cb.global_solve_band(Sb, t1 - t0, bs).double()
        cb.check_status(args[0])
        # dense S (lower) in fp64, damped like both solvers do, for the backward error of each solution
        S = Sd[:n6 * n6].view(n6, n6).double()
        S = torch.tril(S) + torch.tril(S, -1).T
        S = S + torch.diag(torch.diagonal(S) * 1e-4 + 1.0)
        y = Sd[n6 * n6:].double()
        be = lambda x: float((S @ x.reshape(-1) - y).abs().max() / (S.abs().max() * x.abs().max() + y.abs().max()))  # noqa: E731
        # band buffer vs dense matrix, element by element
        ld = (6 * (bs + 2) + 3) // 4 * 4
        r = torch.arange(n6, device=dev)[:, None]; c = torch.arange(n6, device=dev)[None]
        inb = (c <= r) & (r - c < ld)
        band_vals = Sb[(r * ld + c)[inb]]
        dmax = float((band_vals.double() - torch.tril(Sd[:n6 * n6].view(n6, n6)).double()[inb]).abs().max() / S.abs().max())
        rows.append(dict(n=n, M=M, edges=edges, E=int(D["ii"].numel()), span=int(span), bs=int(bs), band_vs_dense_matrix=dmax,
                         dX_maxdiff=float((dXb - dXd).abs().max()), dX_max=float(dXd.abs().max()),
                         backward_err_band=be(dXb), backward_err_dense=be(dXd)))
    return rows


def bench_sharded(dev, rank, world, quick=False, reps=3):
    """BASELINE configs[3] for bench.py: global BA (t0 = 1, all other poses free, 2 GN iterations) patch-sharded over `world` ranks"""
    n, M = (512, 96) if quick else (2048, 192)
    out = {"workload": f"configs[3]: n = {n} keyframes, M = {M} patches/frame, every patch -> frames within +-12 of its source "
                       f"(SURVEY 8d config 4 edge rule, no random loop edges: the reduced system is banded), t0 = 1, 2 GN iterations",
           "n_gpus": world, "scaling": "strong"}
    s = pkg.synthetic.make_ba_scene(seed=5, n=n, M=M, edges="config4", pose_noise=0.005, loop_frac=0.0)
    D = to_dev(s, dev)
    t0, t1 = 1, n
    out["edges"], out["free_poses"] = int(D["ii"].numel()), t1 - t0
    c0 = cost(D, D["poses"], D["patches"])
    times = []
    for i in range(reps + 1):
        ms, poses, patches = run_sharded(D, t0, t1, 2)
        if i >= 1:
            times.append(ms)
    out["ms_2_iterations"] = float(np.median(times))
    out["iters_per_s"] = 2e3 / out["ms_2_iterations"]
    out["pieces_ms_rank0_one_iteration"], out["span_frames"], out["band_half_width"], out["allreduce_bytes_per_iteration"] = pieces(D, t0, t1, rank, world)
    out["solver"] = "band Cholesky (csrc/ba_band.cu)" if out["band_half_width"] else "dense (cuSOLVER potrf via ATen)"
    out["cost_before"], out["cost_after"] = c0, cost(D, poses, patches)
    if world > 1:
        import torch.distributed as dist
        ref = [torch.empty_like(poses) for _ in range(world)]
        dist.all_gather(ref, poses)
        out["ranks_identical"] = bool(all(torch.equal(r, ref[0]) for r in ref))
    elif not quick:
        os.environ["WVB200_GLOBAL_DENSE"] = "1"     # the same problem through the dense path (what round 1 measured)
        try:
            td = [run_sharded(D, t0, t1, 2)[0] for _ in range(2)]
            out["dense_path_ms_2_iterations"] = float(td[-1])
        finally:
            del os.environ["WVB200_GLOBAL_DENSE"]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=512)
    ap.add_argument("--M", type=int, default=96)
    ap.add_argument("--iters", type=int, default=2)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--probe", action="store_true")
    ap.add_argument("--bench", action="store_true", help="print bench.py's `sharded` sub-object")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--edges", default="global", help="global (dpvo.py:448-471 pattern) | config4 | tracker | all")
    ap.add_argument("--loop-frac", type=float, default=0.0, help="config4: share of extra edges to random frames")
    a = ap.parse_args()
    rank, world, local = pkg.dist.init()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _, cb = pkg.native_modules()
    if a.probe:
        for r in probe(dev):
            print(json.dumps(r), flush=True)
        return
    if a.bench:
        out = bench_sharded(dev, rank, world, quick=a.quick)
        if rank == 0:
            print(json.dumps(out), flush=True)
        pkg.dist.finalize()
        return
    s = pkg.synthetic.make_ba_scene(seed=5, n=a.frames, M=a.M, edges=a.edges, pose_noise=0.005, loop_frac=a.loop_frac)
    D = to_dev(s, dev)
    t0, t1 = 1, a.frames
    E = D["ii"].numel()
    times = []
    for i in range(a.reps + 2):
        ms, poses, patches = run_sharded(D, t0, t1, a.iters)
        if i >= 2:
            times.append(ms)
    out = {"what": "global BA, %d GN iterations" % a.iters, "n_gpus": world, "frames": a.frames, "free_poses": t1 - t0, "edges": E,
           "ms": float(np.median(times)), "edges_per_s": E * a.iters / (float(np.median(times)) * 1e-3)}
    out["pieces_ms"], out["span"], out["band_half_width"], out["allreduce_bytes"] = pieces(D, t0, t1, rank, world)
    out["cost_before"], out["cost_after"] = cost(D, D["poses"], D["patches"]), cost(D, poses, patches)
    if a.check:
        _, p1, q1 = run_sharded(D, t0, t1, a.iters, sharded=False)
        out["max_dpose_vs_one_gpu"] = float((poses - p1).abs().max())
        out["max_ddepth_vs_one_gpu"] = float((patches - q1).abs().max())
        out["cost_after_one_gpu"] = cost(D, p1, q1)
        if world > 1:
            import torch.distributed as dist
            ref = [torch.empty_like(poses) for _ in range(world)]
            dist.all_gather(ref, poses)
            out["ranks_identical"] = bool(all(torch.equal(r, ref[0]) for r in ref))
        out["max_pos_err_before"] = float(np.abs(s["poses"][:, :3] - s["poses_gt"][:, :3]).max())
        out["max_pos_err_after"] = float(np.abs(poses.cpu().numpy()[:, :3] - s["poses_gt"][:, :3]).max())
    if rank == 0:
        print(json.dumps(out), flush=True)
    pkg.dist.finalize()


if __name__ == "__main__":
    main()
