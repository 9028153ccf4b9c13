"""Mirror of the reference's COLMAP export, dpvo/plot_utils.py::save_output_for_COLMAP (:58-115): same signature, same three text
files (images.txt, points3D.txt, cameras.txt), byte for byte.  The reference builds the two big files with `str +=` per row
(quadratic in the number of rows); here the rows are formatted in one pass by libwvb200's host-side writer
(csrc/colmap_text.cu: CPython's float repr reimplemented on std::to_chars), the array preparation stays the reference's own
numpy expressions.

`traj` is anything with `.poses_se3` (4x4 world-from-camera matrices) and `.timestamps` -- evo's PoseTrajectory3D as
dpvo_demo.py:189-192 builds it, or `Trajectory(positions_xyz, orientations_quat_wxyz, timestamps)` below for callers without
evo.  evo (==1.27.1, environment.yaml:41) is not part of /root/reference; the two conversions the export needs from it
(xyz + quaternion -> SE3 matrix, SE3 matrix -> quaternion) are restated from its published sources, evo/core/transformations.py
(quaternion_matrix, quaternion_from_matrix with isprecise=False: the eigenvector of the 4x4 K matrix) and evo/core/lie_algebra.py.

nerf_studio_format=True makes the reference shell out to `colmap model_converter` and nerfstudio (plot_utils.py:98-113); neither
exists on this box, so that flag raises NotImplementedError after the text files are written.
"""
import ctypes
import math
import os
from itertools import chain
from pathlib import Path

import numpy as np

from . import capi


# ---- evo restatement ------------------------------------------------------------------------------------------
def quaternion_matrix(quaternion):
    """evo/core/transformations.py::quaternion_matrix (w, x, y, z) -> 4x4"""
    q = np.array(quaternion, dtype=np.float64, copy=True)
    n = np.dot(q, q)
    if n < np.finfo(float).eps * 4.0:
        return np.identity(4)
    q *= math.sqrt(2.0 / n)
    q = np.outer(q, q)
    return np.array([
        [1.0 - q[2, 2] - q[3, 3], q[1, 2] - q[3, 0], q[1, 3] + q[2, 0], 0.0],
        [q[1, 2] + q[3, 0], 1.0 - q[1, 1] - q[3, 3], q[2, 3] - q[1, 0], 0.0],
        [q[1, 3] - q[2, 0], q[2, 3] + q[1, 0], 1.0 - q[1, 1] - q[2, 2], 0.0],
        [0.0, 0.0, 0.0, 1.0]])


def quaternion_from_matrix(matrix):
    """evo/core/transformations.py::quaternion_from_matrix(matrix, isprecise=False) -> (w, x, y, z)"""
    M = np.asarray(matrix, dtype=np.float64)[:4, :4]
    m00, m01, m02 = M[0, 0], M[0, 1], M[0, 2]
    m10, m11, m12 = M[1, 0], M[1, 1], M[1, 2]
    m20, m21, m22 = M[2, 0], M[2, 1], M[2, 2]
    K = np.array([[m00 - m11 - m22, 0.0, 0.0, 0.0],
                  [m01 + m10, m11 - m00 - m22, 0.0, 0.0],
                  [m02 + m20, m12 + m21, m22 - m00 - m11, 0.0],
                  [m21 - m12, m02 - m20, m10 - m01, m00 + m11 + m22]])
    K /= 3.0
    w, V = np.linalg.eigh(K)
    q = V[[3, 0, 1, 2], np.argmax(w)]
    if q[0] < 0.0:
        np.negative(q, q)
    return q


class Trajectory:
    """the slice of evo's PoseTrajectory3D (evo/core/trajectory.py) that the export reads"""

    def __init__(self, positions_xyz=None, orientations_quat_wxyz=None, timestamps=None, poses_se3=None):
        if poses_se3 is None:
            poses_se3 = []
            for xyz, quat in zip(np.array(positions_xyz), np.array(orientations_quat_wxyz)):
                T = np.eye(4)
                T[:3, :3] = quaternion_matrix(quat)[:3, :3]
                T[:3, 3] = xyz
                poses_se3.append(T)
        self.poses_se3 = poses_se3
        self.timestamps = np.array(timestamps)

    @property
    def num_poses(self):
        return len(self.poses_se3)


# ---- the writer ------------------------------------------------------------------------------------------------
def _lib():
    lib = capi()
    lib.wv_colmap_points3d_text.restype = ctypes.c_int64
    lib.wv_colmap_points3d_text.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64]
    lib.wv_colmap_images_text.restype = ctypes.c_int64
    lib.wv_colmap_images_text.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64]
    return lib


def points3d_text(points, colors_uint8, first_id=1):
    """rows of points3D.txt for points [n, 3] (any float dtype; widened to double like .tolist()) and colors [n, 3] uint8"""
    xyz = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
    rgb = np.ascontiguousarray(colors_uint8, dtype=np.uint8).reshape(-1, 3)
    assert len(xyz) == len(rgb)
    buf = np.empty(180 * len(xyz) + 64, np.uint8)
    n = _lib().wv_colmap_points3d_text(xyz.ctypes.data, rgb.ctypes.data, len(xyz), first_id, buf.ctypes.data, buf.size)
    if n < 0:
        raise RuntimeError(f"wv_colmap_points3d_text failed ({n})")
    return buf[:n].tobytes().decode("ascii")


def images_text(quats_wxyz, positions_xyz, names):
    """rows of images.txt (IMAGE_ID, QW, QX, QY, QZ, TX, TY, TZ, CAMERA_ID, NAME + the empty 2D-point line)"""
    pose7 = np.ascontiguousarray(np.concatenate([np.asarray(quats_wxyz, np.float64).reshape(-1, 4),
                                                 np.asarray(positions_xyz, np.float64).reshape(-1, 3)], 1))
    blob = b"".join(n.encode("utf-8") + b"\0" for n in names)
    assert len(names) == len(pose7)
    buf = np.empty(260 * len(pose7) + len(blob) + 64, np.uint8)
    n = _lib().wv_colmap_images_text(pose7.ctypes.data, blob, len(pose7), buf.ctypes.data, buf.size)
    if n < 0:
        raise RuntimeError(f"wv_colmap_images_text failed ({n})")
    return buf[:n].tobytes().decode("utf-8")


def save_output_for_COLMAP(name, tstamp, traj, points, colors, nerf_studio_format, fx, fy, cx, cy, H=480, W=640):
    """Saves the sparse point cloud and camera poses such that it can be opened in COLMAP (dpvo/plot_utils.py:58-115)"""
    colmap_dir = Path(name)
    colmap_dir.mkdir(exist_ok=True)
    scale = 1.0  # for visualization (plot_utils.py:63)

    original_image_path = Path(name).parent.joinpath("images")
    if not os.path.exists(original_image_path):
        original_image_path = Path(name).parent.joinpath("data")
    inv = [np.linalg.inv(p) for p in traj.poses_se3]                      # plot_utils.py:70: camera-from-world
    n = min(len(inv), len(tstamp))                                        # the reference's zip() stops at the shorter one
    positions = np.array([p[:3, 3] for p in inv[:n]]).reshape(-1, 3) * scale
    quats = np.array([quaternion_from_matrix(p) for p in inv[:n]]).reshape(-1, 4)

    image_list = None
    if os.path.exists(original_image_path):
        img_exts = ["*.png", "*.jpeg", "*.jpg"]
        image_list = sorted(chain.from_iterable(Path(original_image_path).glob(e) for e in img_exts))
        if not image_list:
            return
    names = [image_list[int(t)].name if image_list else "image" for t in list(tstamp)[:n]]
    (colmap_dir / "images.txt").write_text(images_text(quats, positions, names))

    colors_uint = (colors * 255).astype(np.uint8)                         # plot_utils.py:90 (the same numpy expression)
    pts = points * scale
    m = min(len(pts), len(colors_uint))
    (colmap_dir / "points3D.txt").write_text(points3d_text(pts[:m], colors_uint[:m]))

    (colmap_dir / "cameras.txt").write_text(f"1 PINHOLE {W} {H} {fx} {fy} {cx} {cy}")

    if nerf_studio_format:
        raise NotImplementedError("nerf_studio_format needs the `colmap` binary and nerfstudio (dpvo/plot_utils.py:98-113): "
                                  "unavailable on this box; the three text files were written")
