"""DPVO.corr as one stacked call at configs[1] (for ncu captures): python tools/corr_one.py"""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
w = pkg.synthetic.make_corr_workload(seed=1)
d = torch.device("cuda")
g = torch.from_numpy(w["gmap"]).half().to(d); f0 = torch.from_numpy(w["fmap0"]).half().to(d); f1 = torch.from_numpy(w["fmap1"]).half().to(d)
c0 = torch.from_numpy(w["coords"]).to(d); ii = torch.from_numpy(w["ii1"]).to(d); jj = torch.from_numpy(w["jj1"]).to(d)
for _ in range(3):
    out = pkg.altcorr.corr_stacked(g, (f0, f1), c0, ii, jj, 3)
torch.cuda.synchronize(); print("ok", tuple(out.shape))
