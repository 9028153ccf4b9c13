#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY -- freezes outputs of the REFERENCE ITSELF as fixtures under tests/golden/.

Runs only in the authoring container (needs /root/reference and oracle/_ref built by build_ref.py).
Every array written is either a seeded input or an output of the reference's own code:
  * ref_cuda_corr / ref_cuda_ba : the reference's .cu/.cpp compiled as host code (build_ref.py)
  * dpvo/ba.py + dpvo/projective_ops.py + dpvo/lietorch/*.py imported unchanged from /root/reference
    over ref_lietorch (their own lietorch_cpu.cpp), with two import shims for absent third-party
    packages: torch_scatter.scatter_sum (-> index_add_) and cv2 (unused on this path).
The fixtures are small (.npz, float32/int64) and are what the `-m "not gpu"` oracle tests and the
`-m gpu` parity tests compare against on the GPU box, where /root/reference does not exist.
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import build_ref  # noqa: E402

syn = importlib.import_module("wild-video-3d-reconstruction_b200.synthetic")
OUT = os.path.join(ROOT, "tests", "golden")
T = torch.from_numpy


def import_reference_python(ref_root="/root/reference"):
    """import dpvo.ba / dpvo.projective_ops from the reference tree, unchanged."""
    sys.modules["lietorch_backends"] = build_ref.load("ref_lietorch")
    sys.modules["cuda_ba"] = build_ref.load("ref_cuda_ba")
    sys.modules["cuda_corr"] = build_ref.load("ref_cuda_corr")
    ts = types.ModuleType("torch_scatter")

    def scatter_sum(src, index, dim=-1, out=None, dim_size=None):
        shape = list(src.shape)
        shape[dim] = dim_size if dim_size is not None else int(index.max()) + 1
        return torch.zeros(shape, dtype=src.dtype).index_add_(dim, index, src)

    ts.scatter_sum = scatter_sum
    sys.modules["torch_scatter"] = ts
    try:
        import cv2  # noqa: F401
    except Exception:
        sys.modules["cv2"] = types.ModuleType("cv2")
    if ref_root not in sys.path:
        sys.path.insert(0, ref_root)
    ba = importlib.import_module("dpvo.ba")
    pops = importlib.import_module("dpvo.projective_ops")
    lie = importlib.import_module("dpvo.lietorch")
    return ba, pops, lie


def corr_cases(corr):
    rng = np.random.default_rng(11)
    out = {}
    for name, (C, radius, dt) in {"f32_r3": (16, 3, torch.float32), "f16_r3": (128, 3, torch.float16),
                                   "f32_r1": (8, 1, torch.float32), "f16_r0": (8, 0, torch.float16)}.items():
        n1, n2, H2, W2, E = 10, 3, 12, 14, 40
        f1 = (0.25 * rng.standard_normal((1, n1, C, 3, 3))).astype(np.float32)
        f2 = (0.25 * rng.standard_normal((1, n2, C, H2, W2))).astype(np.float32)
        co = np.stack([rng.uniform(-6, W2 + 6, (1, E, 3, 3)), rng.uniform(-6, H2 + 6, (1, E, 3, 3))], 2).astype(np.float32)
        co[0, 0] = 5.0  # exact integers (dx = dy = 0)
        co[0, 1] = np.float32(-0.5)
        co[0, 2] = 1e4  # far outside
        ii = rng.integers(0, n1, E).astype(np.int64)
        jj = rng.integers(0, n2, E).astype(np.int64)
        f1t, f2t = T(f1).to(dt), T(f2).to(dt)
        o, = corr.forward(f1t, f2t, T(co), T(ii), T(jj), radius)
        g = rng.standard_normal(o.shape).astype(np.float32)
        g1, g2 = corr.backward(f1t.float(), f2t.float(), T(co), T(ii), T(jj), T(g), radius)
        out.update({f"{name}_fmap1": f1t.float().numpy(), f"{name}_fmap2": f2t.float().numpy(), f"{name}_coords": co,
                    f"{name}_ii": ii, f"{name}_jj": jj, f"{name}_out": o.float().contiguous().numpy(),
                    f"{name}_grad": g, f"{name}_g1": g1.numpy(), f"{name}_g2": g2.numpy()})
    # patchify (correlation_kernel.cu:288-333) at the radii net.py uses (0 and 1)
    net = rng.standard_normal((2, 5, 9, 11)).astype(np.float32)
    pc = np.stack([rng.uniform(-2, 13, (2, 7)), rng.uniform(-2, 11, (2, 7))], -1).astype(np.float32)
    for r in (0, 1, 3):
        p, = corr.patchify_forward(T(net), T(pc), r)
        pg = rng.standard_normal(p.shape).astype(np.float32)
        ng, = corr.patchify_backward(T(net), T(pc), T(pg), r)
        out.update({f"patchify_r{r}_out": p.numpy(), f"patchify_r{r}_grad": pg, f"patchify_r{r}_netgrad": ng.numpy()})
    out.update({"patchify_net": net, "patchify_coords": pc})
    return out


def ba_cases(ba):
    out = {}
    cfgs = {"win": dict(seed=0, n=6, M=8, edges="all"), "trk": dict(seed=3, n=12, M=6, edges="tracker", lifetime=5, removal=8),
            "off": dict(seed=5, n=5, M=4, edges="all", first_frame=3, n_buffer=12)}
    for name, kw in cfgs.items():
        s = syn.make_ba_scene(**kw)
        for tag, (t0, t1, iters) in {"full": (s["t0"], s["t1"], 2), "late": (s["t1"] - 2, s["t1"], 3),
                                     "struct": (s["t1"], s["t1"], 2)}.items():
            poses, patches = T(s["poses"].copy()), T(s["patches"].copy())
            ba.forward(poses, patches, T(s["intrinsics"]), T(s["target"]), T(s["weight"]), T(s["lmbda"]),
                       T(s["ii"]), T(s["jj"]), T(s["kk"]), t0, t1, iters)
            out.update({f"{name}_{tag}_poses": poses.numpy(), f"{name}_{tag}_patches": patches.numpy(),
                        f"{name}_{tag}_args": np.array([t0, t1, iters], np.int64)})
        rp = ba.reproject(T(s["poses"]), T(s["patches"]), T(s["intrinsics"]), T(s["ii"]), T(s["jj"]), T(s["kk"]))
        ix, jx = ba.neighbors(T(s["kk"]), T(s["jj"]))
        out.update({f"{name}_reproject": rp.numpy(), f"{name}_nb_ix": ix.numpy(), f"{name}_nb_jx": jx.numpy()})
        out.update({f"{name}_in_{k}": s[k] for k in ("poses", "patches", "intrinsics", "target", "weight", "lmbda",
                                                      "ii", "jj", "kk", "patches_est")})
    rng = np.random.default_rng(7)
    a, b = rng.integers(0, 9, 300).astype(np.int64), rng.integers(0, 5, 300).astype(np.int64)
    ix, jx = ba.neighbors(T(a), T(b))
    out.update({"nb_rand_ii": a, "nb_rand_jj": b, "nb_rand_ix": ix.numpy(), "nb_rand_jx": jx.numpy()})
    # solve_system (ba.cpp:174-234): chain + one loop edge
    n, r = 6, 7
    ei = np.array([0, 1, 2, 3, 4, 5, 1], np.int64); ej = np.array([1, 2, 3, 4, 5, 0, 4], np.int64)
    Ji = (np.eye(7)[None] + 0.1 * rng.standard_normal((r, 7, 7))).astype(np.float32)
    Jj = (-np.eye(7)[None] + 0.1 * rng.standard_normal((r, 7, 7))).astype(np.float32)
    res = rng.standard_normal((r, 7)).astype(np.float32)
    for tag, freen in {"full": -1, "free4": 4}.items():
        d, = ba.solve_system(T(Ji), T(Jj), T(ei), T(ej), T(res), 1e-4, 1e-4, freen)
        out[f"pgo_{tag}_delta"] = d.numpy()
    out.update({"pgo_Ji": Ji, "pgo_Jj": Jj, "pgo_ii": ei, "pgo_jj": ej, "pgo_res": res})
    return out


def pyba_cases():
    """dpvo/ba.py::BA with the depth prior (:151-159), one GN iteration per call, ep=1.0 so that its damping
    equals ba_cuda.cu:517-518; bounds = the CUDA kernel's fixed +-64 px window (:277-278)."""
    pyba, pops, lie = import_reference_python()
    out = {}
    for name, kw in {"win": dict(seed=0, n=6, M=8, edges="all"), "trk": dict(seed=3, n=12, M=6, edges="tracker", lifetime=5, removal=8)}.items():
        s = syn.make_ba_scene(**kw)
        cx, cy = s["intrinsics"][0, 2], s["intrinsics"][0, 3]
        bounds = [-64, -64, 2 * cx + 64, 2 * cy + 64]
        for tag, est in {"prior": s["patches_est"], "noprior": np.zeros_like(s["patches_est"])}.items():
            G, pat = pyba.BA(lie.SE3(T(s["poses"].copy())[None]), T(s["patches"].copy())[None], T(s["intrinsics"])[None],
                             T(s["target"])[None], T(s["weight"])[None], 1e-4, T(s["ii"]), T(s["jj"]), T(s["kk"]),
                             bounds, ep=1.0, fixedp=s["t0"], patches_est=T(est)[None])
            out.update({f"{name}_{tag}_poses": G.data[0].numpy(), f"{name}_{tag}_patches": pat[0].numpy()})
        co = pops.transform(lie.SE3(T(s["poses"])[None]), T(s["patches"])[None], T(s["intrinsics"])[None],
                            T(s["ii"]), T(s["jj"]), T(s["kk"]))
        out[f"{name}_transform"] = co[0].numpy()
    return out


def pops_cases():
    """dpvo/projective_ops.py (transform / flow_mag / point_cloud) over the reference's own lietorch CPU ops, and the keyframe
    distance + distance-based edge selection of dpvo/dpvo.py:383-429 restated with those same reference functions (the methods
    themselves live on the CUDA-only DPVO class).  Per-frame intrinsics differ from frame to frame on purpose."""
    _, pops, lie = import_reference_python()
    out = {}
    for name, kw in {"trk": dict(seed=31, n=14, M=6, edges="tracker", lifetime=5, removal=8), "glb": dict(seed=32, n=24, M=4, edges="global")}.items():
        s = syn.make_ba_scene(**kw)
        rng = np.random.default_rng(7)
        intr = s["intrinsics"] * (1.0 + 0.05 * rng.standard_normal(s["intrinsics"].shape)).astype(np.float32)
        poses = s["poses"].copy()
        poses[:, 3:] *= (1.0 + 0.01 * rng.standard_normal((len(poses), 1))).astype(np.float32)  # not exactly unit: lietorch renormalises
        patches = s["patches"].copy()
        patches[::9, 2] = 0.01  # far points; a few behind-the-camera cases come from the pose noise
        G = lie.SE3(T(poses)[None]); pt = T(patches)[None]; K = T(intr)[None]
        ii, jj, kk = T(s["ii"]), T(s["jj"]), T(s["kk"])
        out.update({f"{name}_poses": poses, f"{name}_patches": patches, f"{name}_intrinsics": intr,
                    f"{name}_ii": s["ii"], f"{name}_jj": s["jj"], f"{name}_kk": s["kk"]})
        out[f"{name}_transform"] = pops.transform(G, pt, K, ii, jj, kk)[0].numpy()
        co, val = pops.transform(G, pt, K, ii, jj, kk, depth=True, valid=True)
        out[f"{name}_transform_depth"] = co[0].numpy(); out[f"{name}_valid"] = val[0].numpy()
        out[f"{name}_transform_tonly"] = pops.transform(G, pt, K, ii, jj, kk, tonly=True)[0].numpy()
        xj, vj, (Ji, Jj, Jz) = pops.transform(G, pt, K, ii, jj, kk, jacobian=True)       # projective_ops.py:71-99
        out.update({f"{name}_jac_x1": xj[0].numpy(), f"{name}_jac_valid": vj[0].numpy(), f"{name}_jac_Ji": Ji[0].numpy(),
                    f"{name}_jac_Jj": Jj[0].numpy(), f"{name}_jac_Jz": Jz[0].numpy()})
        for beta in (0.3, 0.5):
            out[f"{name}_flow_mag_{beta}"] = pops.flow_mag(G, pt, K, ii, jj, kk, beta=beta)[0].numpy()
        n, M = s["n"], s["M"]
        ix = torch.arange(len(patches)) // M
        ix = ix.clamp(max=len(poses) - 1)
        out[f"{name}_ix"] = ix.numpy()
        out[f"{name}_point_cloud"] = pops.point_cloud(G, pt, K, ix)[0].numpy()
        # compute_keyframe_distance(i, j, beta=0.5), dpvo/dpvo.py:383-408, for all pairs; then get_distance_based_edges, :410-429
        D = np.zeros((n, n), np.float32)
        for i in range(n):
            for j in range(i + 1, n):
                a = torch.full((M,), i, dtype=torch.long); b = torch.full((M,), j, dtype=torch.long)
                ki = torch.arange(M * i, M * (i + 1)); kj = torch.arange(M * j, M * (j + 1))
                fij = pops.flow_mag(G, pt, K, a, b, ki, beta=0.5); fji = pops.flow_mag(G, pt, K, b, a, kj, beta=0.5)
                D[i, j] = D[j, i] = (0.5 * (fij.mean() + fji.mean())).item()
        out[f"{name}_frame_distance"] = D
        # threshold in the widest gap of the middle third of the sorted distances, so that the selection does not hinge on rounding
        dv = np.sort(D[np.triu_indices(n, 2)])
        lo, hi = len(dv) // 3, 2 * len(dv) // 3
        g = int(np.argmax(np.diff(dv[lo:hi]))) + lo
        thresh = float(0.5 * (dv[g] + dv[g + 1]))
        ei, ej = [], []
        for i in range(n - 1):
            ei.append(i); ej.append(i + 1)
        for i in range(n):
            for j in range(i + 2, n):
                if D[i, j] < thresh:
                    ei.append(i); ej.append(j)
        out[f"{name}_thresh"] = np.float32(thresh)
        out[f"{name}_edges_ii"] = np.array(ei, np.int64); out[f"{name}_edges_jj"] = np.array(ej, np.int64)
    return out


def main():
    build_ref.build()
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "corr_ref.npz"), **corr_cases(build_ref.load("ref_cuda_corr")))
    np.savez_compressed(os.path.join(OUT, "ba_ref.npz"), **ba_cases(build_ref.load("ref_cuda_ba")))
    np.savez_compressed(os.path.join(OUT, "pyba_ref.npz"), **pyba_cases())
    np.savez_compressed(os.path.join(OUT, "pops_ref.npz"), **pops_cases())
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()
