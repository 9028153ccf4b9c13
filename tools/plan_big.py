"""edge plan at global-BA size (E = 9.8 M): one global_span call (plan + span).  For ncu launch lists: WVB200_NO_COOP=1 python tools/plan_big.py"""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
_, cb = pkg.native_modules()
s = pkg.synthetic.make_ba_scene(seed=5, n=2048, M=192, edges="config4", pose_noise=0.005, loop_frac=0.0)
D = {k: torch.from_numpy(np.ascontiguousarray(s[k])).cuda() for k in ("poses", "patches", "ii", "jj", "kk")}
for i in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); span = cb.global_span(D["poses"], D["patches"], D["ii"], D["jj"], D["kk"], 1, 2048, True); b.record(); torch.cuda.synchronize()
    print("plan + span", a.elapsed_time(b), "ms, span", span, flush=True)
