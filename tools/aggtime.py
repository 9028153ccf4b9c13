"""SoftAgg at the tracker's size (E = 45 312, D = 384, fp16): the fused plan + kernel next to the same arithmetic composed from torch
ops in the reference's order (unique -> scatter max -> exp -> scatter sum -> div -> mul -> scatter sum; torch_scatter itself is not
in this image).  python tools/aggtime.py"""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
blocks = pkg.blocks
s = pkg.synthetic.make_ba_scene(seed=2, n=40, M=96, edges="tracker")
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, n=10):
    ts = []
    for i in range(n + 3):
        flush.fill_(1)
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(a.elapsed_time(b))
    return float(np.median(ts))
for name, keys in (("kk", s["kk"]), ("ii*12345+jj", s["ii"] * 12345 + s["jj"])):
    ix = torch.from_numpy(keys.astype(np.int64)).cuda()
    E, D = ix.numel(), 384
    f = torch.randn(1, E, D, device="cuda").half(); g = torch.randn(1, E, D, device="cuda").half()
    lo, hi = int(ix.min()), int(ix.max())
    def composed():
        _, jx = torch.unique(ix, return_inverse=True)
        return blocks._soft_agg_autograd(f, g, jx, int(jx.max()) + 1)
    def fused():
        jx, plan = blocks.group_plan(ix, lo, hi)
        return blocks.scatter_softmax_sum(f, g, plan)
    cap = 3840 if name == "kk" else 40 * 40                  # patches in the buffer / frame pairs: what the tracker knows up front
    def fused_nosync():
        jx, plan = blocks.group_plan(ix, lo, hi, g_capacity=cap)
        return blocks.scatter_softmax_sum(f, g, plan)
    jx, plan = blocks.group_plan(ix, lo, hi)
    kernel_only = lambda: blocks.scatter_softmax_sum(f, g, plan)
    if len(sys.argv) > 1 and sys.argv[1] == "kernel":       # sweep mode: the kernel alone
        print(f"{name}: kernel alone {timeit(kernel_only):.4f} ms", flush=True)
        continue
    t_c, t_f, t_k, t_n = timeit(composed), timeit(fused), timeit(kernel_only), timeit(fused_nosync)
    gbs = (2 * E * D * 2 + plan[3] * D * 2) / (t_k * 1e-3) / 1e9
    print(f"{name}: G={plan[3]}  composed torch ops {t_c:.4f} ms   plan+kernel {t_f:.4f} ms (no host read: {t_n:.4f} ms)   kernel alone {t_k:.4f} ms = {gbs:.0f} GB/s algorithmic", flush=True)
