import sys, os, importlib, numpy as np, torch
sys.path.insert(0, "/root/repo")
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
_, cb = pkg.native_modules()
n, M = 1024, 32
s = pkg.synthetic.make_ba_scene(seed=5, n=n, M=M, edges="config4", pose_noise=0.005, loop_frac=0.0)
D = {k: torch.from_numpy(np.ascontiguousarray(s[k])).cuda() for k in ("poses","patches","intrinsics","target","weight","lmbda","ii","jj","kk")}
args = (D["poses"].clone(), D["patches"].clone(), D["intrinsics"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], 1, n)
span = cb.global_span(args[0], args[1], D["ii"], D["jj"], D["kk"], 1, n, True); bs = cb.band_block(span)
Sy = cb.global_build_band(*args, 0, 1, bs)
for _ in range(2):
    dX = cb.global_solve_band(Sy, n - 1, bs)
torch.cuda.synchronize(); print("ok", span, bs, float(dX.abs().max()))
