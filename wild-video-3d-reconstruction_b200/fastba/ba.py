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
