"""BASELINE.json configs[0]: the reference's own PyTorch BA (dpvo/ba.py::BA over its lietorch CPU ops, built as host code in
oracle/_ref) on a synthetic 10-frame window, 96 patches per frame, timed on THIS machine's host cores.  /root/reference does not
exist on the GPU box, so this number can only be taken in the build container; bench.py's reference arm (the reference's fastba /
altcorr kernels compiled as host code) is the one that travels.

    python oracle/time_python_ba.py [iterations]
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import make_golden as mg  # noqa: E402

pyba, pops, lie = mg.import_reference_python()
syn = mg.syn
T = mg.T
for edges in ("all", "tracker"):
    s = syn.make_ba_scene(seed=0, n=10, M=96, edges=edges)
    cx, cy = s["intrinsics"][0, 2], s["intrinsics"][0, 3]
    bounds = [-64, -64, 2 * cx + 64, 2 * cy + 64]
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    ts = []
    for rep in range(4):
        G, pat = lie.SE3(T(s["poses"].copy())[None]), T(s["patches"].copy())[None]
        t = time.perf_counter()
        for _ in range(iters):
            G, pat = pyba.BA(G, pat, T(s["intrinsics"])[None], T(s["target"])[None], T(s["weight"])[None], 1e-4, T(s["ii"]), T(s["jj"]),
                             T(s["kk"]), bounds, ep=1.0, fixedp=s["t0"], patches_est=T(s["patches_est"])[None])
        ts.append(time.perf_counter() - t)
    print(f"dpvo/ba.py::BA, n=10, M=96, edges={edges} (E={len(s['kk'])}), {iters} GN iterations, torch threads={torch.get_num_threads()}: "
          f"{1e3 * np.median(ts[1:]):.1f} ms  ({iters / np.median(ts[1:]):.1f} iterations/s)")
