"""Mirror of the reference's dpvo/projective_ops.py over fused sm_100a kernels (cuda_ba.transform / flow_mag / point_cloud).

Same function names, argument order and result shapes as dpvo/projective_ops.py:19-121, so `from dpvo import projective_ops as
pops` call sites (DPVO.reproject dpvo/dpvo.py:318-321, motionmag / compute_keyframe_distance :360-408, get_pose point clouds)
work unchanged.  `poses` may be a lietorch SE3 (anything with a `.data` tensor [1, N, 7]) or that tensor itself.  One kernel per
call replaces ~12 lietorch / elementwise launches and the repeat()-materialised broadcast of Gij (lietorch/broadcasting.py:26-29).

jacobian=True (used by the Python BA, dpvo/ba.py:95) returns the centre-pixel Jacobians Ji, Jj, Jz from one more kernel.
"""
import torch

from . import native_modules

_, cuda_ba = native_modules()

MIN_DEPTH = 0.2


def _pose_data(poses):
    return poses if torch.is_tensor(poses) else poses.data


def extract_intrinsics(intrinsics):
    return intrinsics[..., None, None, :].unbind(dim=-1)


def coords_grid(ht, wd, **kwargs):
    y, x = torch.meshgrid(torch.arange(ht).to(**kwargs).float(), torch.arange(wd).to(**kwargs).float(), indexing="ij")
    return torch.stack([x, y], dim=-1)


def iproj(patches, intrinsics):
    """inverse projection (projective_ops.py:19-30): [.., 3, P, P] -> [.., P, P, 4]"""
    x, y, d = patches.unbind(dim=2)
    fx, fy, cx, cy = intrinsics[..., None, None].unbind(dim=2)
    return torch.stack([(x - cx) / fx, (y - cy) / fy, torch.ones_like(d), d], dim=-1)


def proj(X, intrinsics, depth=False):
    """projection with d = 1 / Z.clamp(min=0.1) (projective_ops.py:33-49)"""
    X, Y, Z, W = X.unbind(dim=-1)
    fx, fy, cx, cy = intrinsics[..., None, None].unbind(dim=2)
    d = 1.0 / Z.clamp(min=0.1)
    x = fx * (d * X) + cx
    y = fy * (d * Y) + cy
    return torch.stack([x, y, d], dim=-1) if depth else torch.stack([x, y], dim=-1)


def transform(poses, patches, intrinsics, ii, jj, kk, depth=False, valid=False, jacobian=False, tonly=False):
    """projective transform (projective_ops.py:53-104): coords [1, E, P, P, 2|3] (, valid [1, E, P, P])"""
    if jacobian:   # projective_ops.py:71-99: (x1, (Z_centre > 0.2).float(), (Ji, Jj, Jz))
        x1, val, Ji, Jj, Jz = cuda_ba.transform_jacobian(_pose_data(poses), patches, intrinsics, ii, jj, kk, depth, tonly)
        return x1, val, (Ji, Jj, Jz)
    out = cuda_ba.transform(_pose_data(poses), patches, intrinsics, ii, jj, kk, depth, valid, tonly)
    return (out[0], out[1]) if valid else out[0]


def point_cloud(poses, patches, intrinsics, ix):
    """3-D points of all patches in world coordinates (projective_ops.py:106-108): [1, N*M, P, P, 4]"""
    return cuda_ba.point_cloud(_pose_data(poses), patches, intrinsics, ix)


def flow_mag(poses, patches, intrinsics, ii, jj, kk, beta=0.3):
    """beta-blend of full-motion and translation-only flow magnitudes (projective_ops.py:111-121): [1, E, P, P]"""
    return cuda_ba.flow_mag(_pose_data(poses), patches, intrinsics, ii, jj, kk, beta)


def frame_distance(poses, patches, intrinsics, n, M, beta=0.5):
    """D[i, j] = DPVO.compute_keyframe_distance(i, j) (dpvo/dpvo.py:383-408) for all pairs of the first n frames: [n, n]"""
    return cuda_ba.frame_distance(_pose_data(poses), patches, intrinsics, n, M, beta)


def distance_based_edges(poses, patches, intrinsics, n, M, thresh, beta=0.5):
    """DPVO.get_distance_based_edges (dpvo/dpvo.py:410-429) without its O(n^2) host loop: the sequential pairs (i, i+1), then every
    (i, j), j >= i + 2, whose distance is below `thresh`, in the reference's (i-major, j ascending) order.  Returns (ii, jj) lists."""
    if n < 2:
        return [], []
    D = frame_distance(poses, patches, intrinsics, n, M, beta)
    far = torch.triu(D < thresh, diagonal=2).nonzero()  # row-major == the reference's loop order
    seq = torch.arange(n - 1, device=D.device)
    ii = torch.cat([seq, far[:, 0]]).tolist()
    jj = torch.cat([seq + 1, far[:, 1]]).tolist()
    return ii, jj
