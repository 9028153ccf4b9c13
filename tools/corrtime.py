import sys, os, importlib, numpy as np, torch
sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
cc, cb = pkg.native_modules()
syn = pkg.synthetic
w = syn.make_corr_workload(seed=1)
dev = torch.device("cuda")
g = torch.from_numpy(w["gmap"]).half().to(dev)
f0 = torch.from_numpy(w["fmap0"]).half().to(dev); f1 = torch.from_numpy(w["fmap1"]).half().to(dev)
c0 = torch.from_numpy(w["coords"]).to(dev); c1 = (c0 / 4).contiguous()
ii = torch.from_numpy(w["ii1"]).to(dev); jj = torch.from_numpy(w["jj1"]).to(dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
def t(fm, co, n=10):
    ts = []
    for i in range(n + 3):
        flush.fill_(1)
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); cc.forward(g, fm, co, ii, jj, 3); b.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(a.elapsed_time(b))
    return np.median(ts)
# usage: python tools/corrtime.py [generic]      (generic: force the any-shape kernel via WVB200_CORR_TILED=0)
if len(sys.argv) > 1 and sys.argv[1] == "generic":
    os.environ["WVB200_CORR_TILED"] = "0"
print("L0 %.4f ms  L1 %.4f ms" % (t(f0, c0), t(f1, c1)), flush=True)
# the two-level stacked form (DPVO.corr in one call) next to the two-call form + torch.stack
altcorr = importlib.import_module("wild-video-3d-reconstruction_b200.altcorr.correlation")
def t2(fn, n=10):
    ts = []
    for i in range(n + 3):
        flush.fill_(1)
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(a.elapsed_time(b))
    return np.median(ts)
E = ii.numel()
two = lambda: torch.stack([cc.forward(g, f0, c0 / 1, ii, jj, 3)[0], cc.forward(g, f1, c0 / 4, ii, jj, 3)[0]], -1).view(1, E, -1)
one = lambda: altcorr.corr_stacked(g, (f0, f1), c0, ii, jj, 3)
assert torch.equal(two(), one())
print("DPVO.corr: two calls + stack %.4f ms   stacked kernel %.4f ms" % (t2(two), t2(one)), flush=True)
# the generic kernels that carry training shapes (no timing evidence): fp32 forward, backward, patchify fwd / bwd
gf, f0f = g.float(), f0.float()
gr = torch.randn(1, E, 7, 7, 3, 3, device=dev)
print("generic fp32 forward L0 %.4f ms" % t2(lambda: cc.forward(gf, f0f, c0, ii, jj, 3)), flush=True)
print("backward (fp32, L0)     %.4f ms" % t2(lambda: cc.backward(gf, f0f, c0, ii, jj, gr, 3), n=5), flush=True)
net = torch.randn(1, 128, 96, 128, device=dev)
pc = torch.rand(1, 96, 2, device=dev) * torch.tensor([120.0, 90.0], device=dev) + 2
print("patchify fwd r=1        %.4f ms" % t2(lambda: cc.patchify_forward(net, pc, 1)), flush=True)
pg = torch.randn(1, 96, 128, 3, 3, device=dev)
print("patchify bwd r=1        %.4f ms" % t2(lambda: cc.patchify_backward(net, pc, pg, 1)), flush=True)
# the reference's own CUDA build beside it, when oracle/_ref/refgpu_cuda_corr.so was built (oracle/build_ref_gpu.py)
sys.path.insert(0, os.path.join(os.getcwd(), "oracle"))
import build_ref
rc = build_ref.load("refgpu_cuda_corr")
if rc is not None:
    print("REF-CUDA: fp32 forward L0 %.4f ms, backward %.4f ms, fp16 forward L0 %.4f ms" % (
        t2(lambda: rc.forward(gf, f0f, c0, ii, jj, 3), n=5), t2(lambda: rc.backward(gf, f0f, c0, ii, jj, gr, 3), n=5),
        t2(lambda: rc.forward(g, f0, c0, ii, jj, 3), n=5)), flush=True)
    ours = cc.backward(gf, f0f, c0, ii, jj, gr, 3); ref = rc.backward(gf, f0f, c0, ii, jj, gr, 3)
    for a_, b_ in zip(ours, ref):
        print("  backward max |ours - ref| / max |ref| = %.2e" % ((a_ - b_).abs().max().item() / b_.abs().max().item()))
