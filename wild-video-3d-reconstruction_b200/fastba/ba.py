"""Mirror of the reference's dpvo/fastba/ba.py (BA / neighbors / reproject) over this repository's `cuda_ba`."""
from .. import native_modules

_, cuda_ba = native_modules()

neighbors = cuda_ba.neighbors
reproject = cuda_ba.reproject
solve_system = cuda_ba.solve_system


def BA(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, iterations=2, patches_est=None, mu=2.0):
    """In-place Gauss-Newton BA (dpvo/fastba/ba.py:7-8).  `patches_est`/`mu` switch on the depth-prior term of
    dpvo/ba.py:151-159; leaving them out is the reference's 12-argument behaviour."""
    return cuda_ba.forward(poses.data, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, iterations,
                           patches_est, mu)


def BA_sharded(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, iterations=2, patches_est=None, mu=2.0,
               group=None):
    """fastba.BA for the global problem, patch-sharded over the ranks of `group`.

    Every rank holds the full state (poses, patches, edge list) and calls this collectively.  Per Gauss-Newton iteration rank r
    accumulates the partial reduced system [S | y] of ITS contiguous run of patches (cuda_ba.global_build), ONE all_reduce sums
    the flat buffer, every rank runs the same solve (band Cholesky, csrc/ba_band.cu, when the edge list makes S banded; else the dense
    cuda_ba.global_solve: damping + Cholesky, ba_cuda.cu:517-523), rank r
    back-substitutes its patches' depth updates (cuda_ba.global_dz), a second all_reduce sums the dZ vector, and every rank
    applies the same retraction -- so the replicas of the state stay identical without a broadcast.
    Without an initialised process group this is the single-GPU global path."""
    import torch.distributed as dist

    world, rank = 1, 0
    if dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    poses = poses.data
    E = ii.numel()
    if E == 0 or iterations <= 0:
        return
    # Banded or dense?  The widest frame span of any patch's observations decides (one host sync; it also builds the edge plan that
    # every piece below reuses).  Every rank holds the same edge list, so every rank takes the same branch.
    bs = cuda_ba.band_block(cuda_ba.global_span(poses, patches, ii, jj, kk, t0, t1, True)) if t1 - t0 > 0 else 0
    for it in range(iterations):
        if bs:
            # band buffer [band | y]: 6N * (6 (bs + 2) + 1) floats -- 15 MB at N = 2047, bs = 24 -- instead of the dense 603 MB
            Sy = cuda_ba.global_build_band(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, rank, world, bs,
                                           patches_est, mu)
        else:
            Sy = cuda_ba.global_build(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, rank, world, False,
                                      patches_est, mu)
        cuda_ba.check_status(poses)  # bad index / over-dense patch: say so before the solve trips over it
        if world > 1:
            dist.all_reduce(Sy, group=group)
        dX = cuda_ba.global_solve_band(Sy, t1 - t0, bs) if bs else cuda_ba.global_solve(Sy, t1 - t0)
        del Sy
        dZ = cuda_ba.global_dz(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, rank, world, dX, patches_est, mu)
        if world > 1:
            dist.all_reduce(dZ, group=group)
        cuda_ba.global_retract(poses, patches, E, t0, t1, dX, dZ)
    cuda_ba.check_status(poses)  # bad index / over-dense patch / non-SPD pivot flagged by any piece above (one host sync, like cuda_ba.forward)
