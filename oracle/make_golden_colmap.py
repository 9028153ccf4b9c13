"""Golden vectors for the COLMAP text export: runs the reference's OWN save_output_for_COLMAP (dpvo/plot_utils.py:58-115, loaded
from /root/reference by file path) on seeded inputs and commits its three text files + the inputs under tests/golden/colmap_ref/.

The module imports evo, matplotlib, loguru and formatter.colmap_utilis at import time; none of them is installed here, so they are
stubbed: the plotting / logging / nerfstudio stubs are never reached by the export, and evo's PoseTrajectory3D is a restatement of
evo==1.27.1's class (evo/core/trajectory.py: lazily derived positions_xyz / orientations_quat_wxyz / poses_se3) over the
conversion functions restated in wild-video-3d-reconstruction_b200/colmap_export.py.  So the TEXT layer (row layout, number
formatting, colour conversion, image-name lookup) is pinned by the reference's code itself; evo's matrix <-> quaternion arithmetic
is a restatement (parity unpinned for that step).

    python oracle/make_golden_colmap.py
"""
import importlib
import importlib.util
import os
import shutil
import sys
import tempfile
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ce = importlib.import_module("wild-video-3d-reconstruction_b200.colmap_export")


class PoseTrajectory3D:
    def __init__(self, positions_xyz=None, orientations_quat_wxyz=None, timestamps=None, poses_se3=None, meta=None):
        if positions_xyz is not None:
            self._positions_xyz = np.array(positions_xyz)
        if orientations_quat_wxyz is not None:
            self._orientations_quat_wxyz = np.array(orientations_quat_wxyz)
        if poses_se3 is not None:
            self._poses_se3 = poses_se3
        self.timestamps = np.array(timestamps)

    @property
    def positions_xyz(self):
        if not hasattr(self, "_positions_xyz"):
            self._positions_xyz = np.array([p[:3, 3] for p in self._poses_se3])
        return self._positions_xyz

    @property
    def orientations_quat_wxyz(self):
        if not hasattr(self, "_orientations_quat_wxyz"):
            self._orientations_quat_wxyz = np.array([ce.quaternion_from_matrix(p) for p in self._poses_se3])
        return self._orientations_quat_wxyz

    @property
    def poses_se3(self):
        if not hasattr(self, "_poses_se3"):
            self._poses_se3 = ce.Trajectory(self.positions_xyz, self.orientations_quat_wxyz, self.timestamps).poses_se3
        return self._poses_se3

    @property
    def num_poses(self):
        return len(self._poses_se3) if hasattr(self, "_poses_se3") else self.positions_xyz.shape[0]


def load_reference_plot_utils():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m
    mod("formatter"); mod("formatter.colmap_utilis", colmap_to_json=lambda *a, **k: None)
    mod("matplotlib"); mod("matplotlib.pyplot")
    mod("evo"); mod("evo.core"); mod("evo.core.sync"); mod("evo.tools"); mod("evo.tools.plot")
    mod("evo.core.trajectory", PoseTrajectory3D=PoseTrajectory3D)
    sys.modules["evo.core"].sync = sys.modules["evo.core.sync"]
    sys.modules["evo.tools"].plot = sys.modules["evo.tools.plot"]
    log = types.SimpleNamespace(info=lambda *a, **k: None, error=lambda *a, **k: None)
    mod("loguru", logger=log)
    spec = importlib.util.spec_from_file_location("ref_plot_utils", "/root/reference/dpvo/plot_utils.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def make_inputs(seed, n_poses, n_points):
    rng = np.random.default_rng(seed)
    q = rng.standard_normal((n_poses, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    poses = np.concatenate([rng.uniform(-3, 3, (n_poses, 3)), q], 1).astype(np.float32)       # tx ty tz qx qy qz qw (dpvo_demo.py:188)
    tstamps = np.arange(n_poses, dtype=np.float64)
    points = (rng.standard_normal((n_points, 3)) * np.array([5, 2, 20])).astype(np.float32)
    points[:6] = np.array([[0, -0.0, 1e-5], [1e16, 123456.789, 1e-7], [0.1, 0.25, 3], [1e22, -1e-4, 9.999999e15], [65504, 1e17, -5e-5],
                           [3.4e38, 1.17549435e-38, 1e-45]], np.float32)
    colors = (rng.uniform(0, 1, (n_points, 3)) * 255.0).astype(np.float32)                    # already x255 (dpvo/dpvo.py:259)
    calib = tuple(np.array([fx, fy, cx, cy], np.float32) for fx, fy, cx, cy in [(440.0, 441.5, 256.25, 192.0)])[0]
    return poses, tstamps, points, colors, calib


def main():
    ref = load_reference_plot_utils()
    out_dir = os.path.join(ROOT, "tests", "golden", "colmap_ref")
    shutil.rmtree(out_dir, ignore_errors=True)
    os.makedirs(out_dir)
    for case, (seed, n_poses, n_points, with_images) in {"plain": (1, 12, 300, False), "named": (2, 7, 40, True)}.items():
        poses, tstamps, points, colors, calib = make_inputs(seed, n_poses, n_points)
        traj = PoseTrajectory3D(positions_xyz=poses[:, :3], orientations_quat_wxyz=poses[:, [6, 3, 4, 5]], timestamps=tstamps)
        with tempfile.TemporaryDirectory() as td:
            if with_images:
                os.makedirs(os.path.join(td, "images"))
                for i in range(n_poses):
                    open(os.path.join(td, "images", f"frame_{i:05d}.{'png' if i % 2 else 'jpg'}"), "w").close()
            cwd = os.getcwd()
            ref.save_output_for_COLMAP(os.path.join(td, "colmap"), tstamps, traj, points, colors, False, *calib, 384, 512)
            os.chdir(cwd)
            texts = {f: open(os.path.join(td, "colmap", f)).read() for f in ("images.txt", "points3D.txt", "cameras.txt")}
        np.savez_compressed(os.path.join(out_dir, case + ".npz"), poses=poses, tstamps=tstamps, points=points, colors=colors, calib=calib,
                            with_images=with_images, **{k.replace(".", "_"): np.frombuffer(v.encode(), np.uint8) for k, v in texts.items()})
        print(case, {k: len(v) for k, v in texts.items()})


if __name__ == "__main__":
    main()
