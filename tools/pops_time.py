"""Device times of the fused projective ops: transform / flow_mag at the tracker's edge count, frame_distance at config-4 size."""
import importlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
pops = pkg.projective_ops


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


s = pkg.synthetic.make_ba_scene(seed=2, n=40, M=96, edges="tracker")
a = [torch.from_numpy(np.ascontiguousarray(s[k])).cuda() for k in ("poses", "patches", "intrinsics", "ii", "jj", "kk")]
out = {"E": int(a[3].numel()),
       "transform_ms": timed(lambda: pops.transform(*a)),
       "flow_mag_ms": timed(lambda: pops.flow_mag(*a, beta=0.5))}
for n, M in ((512, 96), (2048, 192)):
    g = pkg.synthetic.make_ba_scene(seed=5, n=n, M=M, edges="global", pose_noise=0.005)
    b = [torch.from_numpy(np.ascontiguousarray(g[k])).cuda() for k in ("poses", "patches", "intrinsics")]
    ms = timed(lambda: pops.frame_distance(*b, n, M, beta=0.5), reps=3)
    D = pops.frame_distance(*b, n, M, beta=0.5)
    ii, jj = pops.distance_based_edges(*b, n, M, float(D[D > 0].median()) * 0.1, beta=0.5)
    out[f"frame_distance_n{n}_M{M}_ms"] = ms
    out[f"pairs_n{n}"] = n * (n - 1) // 2
    out[f"edges_selected_n{n}"] = len(ii)
print(json.dumps(out))
