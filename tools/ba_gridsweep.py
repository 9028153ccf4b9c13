"""fastba.BA (windowed, 10 free poses, 2 GN iterations) against the number of CTAs of its cooperative grid (WVB200_BA_GRID), at the
tracker's edge counts.  python tools/ba_gridsweep.py"""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
_, cb = pkg.native_modules()
for M in (96, 384):
    s = pkg.synthetic.make_ba_scene(seed=2, n=40, M=M, edges="tracker")
    D = {k: torch.from_numpy(np.ascontiguousarray(s[k])).cuda() for k in ("poses", "patches", "intrinsics", "target", "weight", "lmbda", "ii", "jj", "kk")}
    t1 = int(s["t1"]); t0 = t1 - 10
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for grid in (0, 148, 128, 111, 96, 74, 64, 48, 37, 32, 24, 16):
        if grid: os.environ["WVB200_BA_GRID"] = str(grid)
        ts = []
        for i in range(30):
            po, pa = D["poses"].clone(), D["patches"].clone()
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            cb.forward(po, pa, D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, 2)
            b.record(); torch.cuda.synchronize()
            if i >= 5: ts.append(a.elapsed_time(b))
        print(f"M={M} E={len(s['kk'])} grid={grid or 'default'}: {np.median(ts)*1e3:.1f} us (min {np.min(ts)*1e3:.1f})", flush=True)
