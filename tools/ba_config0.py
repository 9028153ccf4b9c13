"""BASELINE.json configs[0] on the B200: fastba.BA (this build) on the synthetic 10-frame window, 96 patches per frame, that
oracle/time_python_ba.py runs through the reference's PyTorch BA (dpvo/ba.py) on host cores.  python tools/ba_config0.py"""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
_, cb = pkg.native_modules()
for edges in ("all", "tracker"):
    s = pkg.synthetic.make_ba_scene(seed=0, n=10, M=96, edges=edges)
    D = {k: torch.from_numpy(np.ascontiguousarray(s[k])).cuda() for k in ("poses", "patches", "intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk", "patches_est")}
    t0, t1 = int(s["t0"]), int(s["t1"])
    ts = []
    for i in range(23):
        po, pa = D["poses"].clone(), D["patches"].clone()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        cb.forward(po, pa, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, 2, D["patches_est"], 2.0)
        b.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(a.elapsed_time(b))
    print(f"fastba.BA (B200), n=10, M=96, edges={edges} (E={len(s['kk'])}), free poses {t1 - t0}, 2 GN iterations with the depth prior: "
          f"{np.median(ts):.4f} ms ({2e3 / np.median(ts):.0f} iterations/s)", flush=True)
