#!/usr/bin/env python
"""bench.py -- one JSON line for the altcorr + fastba hot path (BASELINE.json configs[1] sizes).

A "step" is what the unchanged tracker does once per frame inside DPVO.update() on the path this repo
replaces (dpvo/dpvo.py:711-749): DPVO.corr (dpvo/dpvo.py:325-333: altcorr.corr on both pyramid levels, E = 45 312 edges, fp16
maps, radius 3, stacked into [1, E, 882]), fastba.neighbors, and fastba.BA with 2 Gauss-Newton iterations on the same edge list
(10 free poses).
One step == one frame, so the metric is frames/s of the hot path at 512x384.

  python bench.py [--gpus N] [--steps K] [--warmup W]             our arm (CUDA, through the torch extensions -> C ABI)
  python bench.py --impl reference ...                            the reference's own code on the host cores

Multi-GPU: the path does not shard inside a clip (replicas only); rank r tracks its own synthetic
clip (seed offset r), no data-path collective, "scaling": "weak".
"""
import argparse
import ctypes
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = "wild-video-3d-reconstruction_b200"
for p in (ROOT,):
    if p not in sys.path:
        sys.path.insert(0, p)

# algorithmic bytes per edge and level for fp16 maps (gmap patch + window footprint + coords +
# indices + output):  s*C*P^2 + s*C*F + 4*2*P^2 + 2*8 + s*49*P^2,  F0 = 100 px, F1 = 81 px
CORR_BYTES_L0 = 2 * 128 * 9 + 2 * 128 * 100 + 72 + 16 + 2 * 49 * 9   # 28 874
CORR_BYTES_L1 = 2 * 128 * 9 + 2 * 128 * 81 + 72 + 16 + 2 * 49 * 9    # 24 010
# dram__bytes_read.sum + dram__bytes_write.sum of one DPVO.corr call = restage_kernel x2 + corr_tc5_kernel<2>, ncu --set full
# (profiles/r1_stacked_ncu_summary.txt)
NCU_TRAFFIC_CORR = 643.0e6
WORKLOAD = ("configs[1]: 512x384 -> fmaps 96x128 + 24x32, C=128 fp16, pmem=36, M=96, tracker edge list E=45312, "
            "radius 3; one update(): DPVO.corr (both levels, stacked [1,E,882]) + neighbors + BA(10 free poses, 2 GN iterations)")


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def make_inputs(seed_offset=0, small=False):
    """numpy inputs of one step (seeded; rank r uses its own clip).  small=True: a reduced copy for the CPU tests."""
    syn = importlib.import_module(PKG + ".synthetic")
    if small:
        w = syn.make_corr_workload(seed=1 + 1000 * seed_offset, M=4, pmem=8, n_frames=10, ht=24, wd=32, lifetime=4, removal=6)
        s = syn.make_ba_scene(seed=2 + 1000 * seed_offset, n=10, M=4, edges="tracker", lifetime=4, removal=6)
        return w, s
    w = syn.make_corr_workload(seed=1 + 1000 * seed_offset)
    s = syn.make_ba_scene(seed=2 + 1000 * seed_offset, n=40, M=96, edges="tracker")
    assert w["E"] == len(s["kk"]) == 45312
    return w, s


class ClockSampler:
    """SM clock + clock-event (throttle) reasons sampled DURING the timed region.

    The timed region of this workload is a few tens of milliseconds, shorter than one `nvidia-smi` start-up, so the samples come
    from NVML in a thread (the same counters `nvidia-smi --query-gpu=clocks.sm,clocks_event_reasons.*` prints, B200_PROFILING.md),
    one every ~1 ms; `nvidia-smi -lms` is only the fallback when NVML cannot be loaded."""
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, torch_index):
        import threading
        self.sm, self.mx, self.reasons = [], None, set()
        self.stop_flag = threading.Event()
        self.thread = None
        self.h = None
        self.smi = None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            self.nv = pynvml
            pr = torch.cuda.get_device_properties(torch_index)
            try:
                bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
                self.h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(torch_index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None
        self.index = torch_index

    def _loop(self):
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for name, bit in self.BITS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self.stop_flag.wait(0.001)

    def start(self):
        import threading
        if self.h is not None:
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()
            return
        try:  # fallback
            self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            self.smi = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                         "-lms", "20"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.smi = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": self.mx, "reasons": [], "source": "nvml thread, 1 ms period"}
        if self.thread is not None:
            self.stop_flag.set()
            self.thread.join(timeout=2)
        elif self.smi is not None:
            out["source"] = "nvidia-smi -lms 20"
            self.smi.terminate()
            try:
                self.smi.wait(timeout=5)
            except Exception:
                self.smi.kill()
            self.f.flush(); self.f.seek(0)
            for line in self.f.read().splitlines():
                c = [x.strip() for x in line.split(",")]
                try:
                    self.sm.append(float(c[0])); self.mx = max(self.mx or 0.0, float(c[1]))
                except (ValueError, IndexError):
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(name)
            os.unlink(self.f.name)
            out["sm_max_mhz"] = self.mx
        if self.sm:
            out["sm_mhz"] = float(np.median(self.sm))
        out["reasons"] = sorted(self.reasons)
        out["samples"] = len(self.sm)
        return out


# ----------------------------------------------------------------------------------------------------------
# reference / CPU legs.  These are the ONLY places bench.py executes anything under oracle/.
# ----------------------------------------------------------------------------------------------------------
def _ref_corr_worker(args):
    """one process: the reference's corr_forward (its own correlation_kernel.cu built as host code) on a slice of edges"""
    lo, hi = args
    import torch
    torch.set_num_threads(1)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import build_ref
    rc = build_ref.load("ref_cuda_corr")
    g = _REF_SHARED
    t = time.perf_counter()
    cs = [rc.forward(g["gmap"], fm, (g["coords"][:, lo:hi] / div).contiguous(), g["ii"][lo:hi], g["jj"][lo:hi], 3)[0].permute(0, 1, 3, 2, 4, 5)
          for fm, div in ((g["f0"], 1.0), (g["f1"], 4.0))]
    torch.stack(cs, -1).view(1, hi - lo, -1)   # dpvo/altcorr/correlation.py:30 + dpvo/dpvo.py:333
    return time.perf_counter() - t


_REF_SHARED = None


def reference_step_seconds(w, s, sample_edges, cores):
    """seconds the reference's own code takes for ONE step on `cores` host cores, extrapolated from a bounded sample:
    corr on `sample_edges` of the E edges (both levels, edges split over `cores` processes; corr cost is linear in the
    edge count), BA + neighbors on the full edge list.  Returns (t_step, t_corr_full, t_ba, kind, sample)."""
    global _REF_SHARED
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import build_ref
    E = w["E"]
    have_ref = all(os.path.exists(os.path.join(ROOT, "oracle", "_ref", n + ".so")) for n in ("ref_cuda_corr", "ref_cuda_ba"))
    Es = min(E, sample_edges)
    if have_ref:
        import multiprocessing as mp
        kind = "reference"
        _REF_SHARED = dict(gmap=torch.from_numpy(w["gmap"]).half(), f0=torch.from_numpy(w["fmap0"]).half(),
                           f1=torch.from_numpy(w["fmap1"]).half(), coords=torch.from_numpy(w["coords"]),
                           ii=torch.from_numpy(w["ii1"]), jj=torch.from_numpy(w["jj1"]))
        per = (Es + cores - 1) // cores
        slices = [(a, min(Es, a + per)) for a in range(0, Es, per)]
        ctx = mp.get_context("fork")
        t = time.perf_counter()
        with ctx.Pool(len(slices)) as pool:
            pool.map(_ref_corr_worker, slices)
        t_corr = (time.perf_counter() - t) * (E / Es)
        rb = build_ref.load("ref_cuda_ba")
        torch.set_num_threads(cores)
        T = lambda k: torch.from_numpy(s[k].copy())  # noqa: E731
        po, pa = T("poses"), T("patches")
        t = time.perf_counter()
        rb.neighbors(T("kk"), T("jj"))
        rb.forward(po, pa, T("intrinsics"), T("target"), T("weight"), T("lmbda"), T("ii"), T("jj"), T("kk"), s["t1"] - 10, s["t1"], 2)
        t_ba = time.perf_counter() - t
    else:
        import oracle as orc
        kind = "port"
        g = torch.from_numpy(w["gmap"]).half().float().numpy()
        t = time.perf_counter()
        for fm, div in ((w["fmap0"], 1.0), (w["fmap1"], 4.0)):
            f = torch.from_numpy(fm).half().float().numpy()
            orc.corr_forward(g, f, (w["coords"][:, :Es] / div).astype(np.float32), w["ii1"][:Es], w["jj1"][:Es], 3, mode=orc.F16EMU, nthreads=cores)
        t_corr = (time.perf_counter() - t) * (E / Es)
        t = time.perf_counter()
        orc.neighbors(s["kk"], s["jj"])
        orc.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"], s["kk"],
                       s["t1"] - 10, s["t1"], 2, mode=orc.F32)
        t_ba = time.perf_counter() - t
    sample = (f"corr: {Es} of {E} edges x 2 levels over {cores} processes, scaled by E/{Es}; neighbors + BA(2 iters): full edge list; "
              + ("reference's own correlation_kernel.cu / ba_cuda.cu / ba.cpp compiled as host code (oracle/_ref)" if have_ref
                 else "oracle port (oracle/liboracle.so), fp16-emulating corr"))
    return t_corr + t_ba, t_corr, t_ba, kind, sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    w, s = make_inputs(0)
    sample_edges = max(cores * 16, 4096)
    times = []
    for it in range(args.warmup + args.steps):
        t_step, t_corr, t_ba, kind, sample = reference_step_seconds(w, s, sample_edges, cores)
        if it >= args.warmup:
            times.append((t_step, t_corr, t_ba))
    t_step = float(np.mean([t[0] for t in times]))
    fps = args.gpus / t_step  # N replicas of the CPU path would need N hosts; reported per the same weak-scaling rule
    line = {
        "impl": "reference", "metric": "dpvo_hotpath_frames_per_s", "value": 1.0 / t_step, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f16 maps / f32 accumulate+geometry", "data": "synthetic",
        "config": {"workload": WORKLOAD},
        "cpu_baseline": {"value": 1.0 / t_step, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample,
                         "corr_s_per_step": float(np.mean([t[1] for t in times])), "ba_s_per_step": float(np.mean([t[2] for t in times]))},
        "e2e": {"value": 1.0 / t_step, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    del fps
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------------------
def run_b200(args):
    # keep stdout for the ONE JSON line: libraries (NCCL's version banner) print there too
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py (our arm) needs a GPU: there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    pkg = importlib.import_module(PKG)
    cc, cb = pkg.native_modules()
    altcorr = importlib.import_module(PKG + ".altcorr.correlation")
    lib = pkg.capi()
    assert lib.wv_device_supported() == 1, "libwvb200 has sm_100a code only"

    w, s = make_inputs(rank)
    E = w["E"]
    pin = lambda a, dt=None: (torch.from_numpy(np.ascontiguousarray(a)).to(dt) if dt else torch.from_numpy(np.ascontiguousarray(a))).pin_memory()  # noqa: E731
    H = dict(gmap=pin(w["gmap"], torch.half), f0=pin(w["fmap0"], torch.half), f1=pin(w["fmap1"], torch.half),
             c0=pin(w["coords"]), c1=pin((w["coords"] / 4.0).astype(np.float32)), ii1=pin(w["ii1"]), jj1=pin(w["jj1"]),
             poses=pin(s["poses"]), patches=pin(s["patches"]), intr=pin(s["intrinsics"]), target=pin(s["target"]),
             weight=pin(s["weight"]), lmbda=pin(s["lmbda"]), ii=pin(s["ii"]), jj=pin(s["jj"]), kk=pin(s["kk"]))
    D = {k: v.to(dev) for k, v in H.items()}
    t0, t1 = int(s["t1"]) - 10, int(s["t1"])
    poses, patches = D["poses"].clone(), D["patches"].clone()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    marks = {}

    def step(timed):
        # state reset + L2 flush are outside the timed events
        poses.copy_(D["poses"]); patches.copy_(D["patches"])
        flush.fill_(1)
        e = [ev() for _ in range(4)]
        e[0].record()
        corr = altcorr.corr_stacked(D["gmap"], (D["f0"], D["f1"]), D["c0"], D["ii1"], D["jj1"], 3)   # DPVO.corr, one call
        e[1].record()
        ix, jx = cb.neighbors(D["kk"], D["jj"])
        e[2].record()
        cb.forward(poses, patches, D["intr"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, 2)
        e[3].record()
        if timed:
            marks.setdefault("ev", []).append(e)
        return corr, ix

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step(False)
    barrier()
    sampler = ClockSampler(local)  # every rank samples its own GPU; rank 0 reports its own line and the slowest clock seen
    sampler.start()
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        step(True)
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop()
    seg = np.array([[a.elapsed_time(b) for a, b in zip(e[:-1], e[1:])] for e in marks["ev"]])  # ms: corr (both levels, stacked), neighbors, BA
    t_dev = float(seg.sum())  # ms over the K timed steps, flush/reset excluded
    tt = torch.tensor([t_dev], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_max = float(tt.item())
    # per-rank segment times and clocks, for the record (the value above is the max over ranks)
    per_rank = None
    if world > 1:
        mine = torch.tensor([float(seg[:, 0].mean()), 0.0, float(seg[:, 1].mean()), float(seg[:, 2].mean())] + [clocks["sm_mhz"] or 0.0, float(len(clocks["reasons"]))], device=dev,
                            dtype=torch.float64)
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = [[round(float(v), 4) for v in r.tolist()] for r in allr]
        clocks["sm_mhz_min_over_ranks"] = min(r[4] for r in per_rank)

    # ---- end to end through the C ABI with HOST buffers (pinned), copies inside the timed region -----------------
    # the call a non-torch embedder makes: hostapi.Session.update == wv_session_update_host (include/wvb200.h)
    T = 7
    # per-level outputs here: the coarse level's volume goes back to the host while the level-0 ring is still being uploaded
    # (the stacked layout interleaves the levels, so its download can only start after everything: 231 vs 286 frames/s)
    out0 = torch.empty((1, E, T, T, 3, 3), dtype=torch.half).pin_memory()
    out1 = torch.empty_like(out0).pin_memory()
    hp, hq = H["poses"].clone().pin_memory(), H["patches"].clone().pin_memory()
    nbx, nbj = torch.empty(E, dtype=torch.long).pin_memory(), torch.empty(E, dtype=torch.long).pin_memory()
    sess = pkg.hostapi.Session()
    ba = dict(poses=hp, patches=hq, intrinsics=H["intr"], target=H["target"], weight=H["weight"], lmbda=H["lmbda"],
              ii=H["ii"], jj=H["jj"], kk=H["kk"], t0=t0, t1=t1, nb_ix=nbx, nb_jx=nbj)
    uargs = sess.make_args(H["gmap"], [H["f0"], H["f1"]], H["c0"], H["ii1"], H["jj1"], [out0, out1], radius=3, ba=ba, iterations=2)

    def e2e_step():
        hp.copy_(H["poses"]); hq.copy_(H["patches"])
        sess.run(uargs)

    nb = lambda *ks: sum(H[k].numel() * H[k].element_size() for k in ks)  # noqa: E731
    h2d = nb("gmap", "f0", "f1", "c0", "ii1", "jj1", "poses", "patches", "intr", "target", "weight", "lmbda", "ii", "jj", "kk")
    d2h = 2 * out0.numel() * 2 + nb("poses", "patches") + 2 * E * 8 + 4
    for _ in range(3):
        e2e_step()
    # the e2e result is the product's result: check it once against the device-resident path of the timed loop above
    cd, ixd = step(False)
    torch.cuda.synchronize()
    cdl = cd.view(1, E, T, T, 3, 3, 2)
    assert torch.equal(out0, cdl[..., 0].cpu()) and torch.equal(out1, cdl[..., 1].cpu()) and torch.equal(nbx, ixd.cpu()), "e2e != device path"
    # ... and the stacked volume is the two per-level calls the reference makes, stacked (dpvo/dpvo.py:331-333)
    two = torch.stack([cc.forward(D["gmap"], D["f0"], D["c0"], D["ii1"], D["jj1"], 3)[0],
                       cc.forward(D["gmap"], D["f1"], D["c1"], D["ii1"], D["jj1"], 3)[0]], -1).view(1, E, -1)
    assert torch.equal(cd, two), "stacked corr != two-call form"
    del two
    assert torch.equal(hp, poses.cpu()), "e2e BA != device path"
    barrier()
    te = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    te = time.perf_counter() - te
    tte = torch.tensor([te], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tte, op=dist.ReduceOp.MAX)
    te = float(tte.item())
    sess.close()

    if rank == 0:
        K = args.steps
        peak, peak_src = measured_peak_gbs()
        corr_ms = float(seg[:, 0].mean())
        achieved = (CORR_BYTES_L0 + CORR_BYTES_L1) * E / (corr_ms * 1e-3) / 1e9
        line = {
            "metric": "dpvo_hotpath_frames_per_s", "value": world * K / (t_max * 1e-3), "unit": "frames/s", "n_gpus": world, "steps": K,
            "warmup": max(args.warmup, 3), "ms_per_step": t_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16 maps / f32 accumulate+geometry", "data": "synthetic",
            "config": {"workload": WORKLOAD, "parallelism": f"replicas x{world} (one clip per GPU, no collective)",
                       "l2": "512 MB fill before every timed step (outside the CUDA-event intervals); maps+outputs are 210 MB > L2"},
            "ms_breakdown": {"corr_both_levels_stacked": corr_ms, "neighbors": float(seg[:, 1].mean()),
                             "ba_2iters": float(seg[:, 2].mean())},
            "corr_gbs": achieved,
            "ba_iters_per_s": 2.0 / (float(seg[:, 2].mean()) * 1e-3),
            "wall_ms_per_step_incl_flush": 1e3 * wall / K,
            "roofline": {"kernel": "DPVO.corr: corr forward on both levels, stacked (restage_kernel x2 + corr_tc5_kernel<2>, fp16 maps)",
                         "bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": NCU_TRAFFIC_CORR, "peak_source": peak_src,
                         "traffic_source": "dram__bytes_read+write of restage_kernel x2 + corr_tc5_kernel<2>, ncu --set full, profiles/r1_stacked_ncu_summary.txt",
                         "algorithmic_bytes_per_launch": (CORR_BYTES_L0 + CORR_BYTES_L1) * E,
                         "dram_gbs_from_traffic": NCU_TRAFFIC_CORR / (corr_ms * 1e-3) / 1e9,
                         "note": "frac > 1: the SURVEY 8(d) per-edge bytes count every edge's window as an HBM read; neighbouring "
                                 "windows overlap and the L2 serves them (hit rate 76 %), so DRAM moves traffic/achieved of it. "
                                 "The tile kernel's own ceiling is the L2->SM tile stream: 90.6 k items x 28 KB per call "
                                 "(DESIGN.md 3.7)"},
            "e2e": {"value": world * K / te, "unit": "frames/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "api": "hostapi.Session.update -> wv_session_update_host (C ABI, pinned host buffers, 3 streams)"},
            "per_rank": {"columns": ["corr_ms", "-", "neighbors_ms", "ba_ms", "sm_mhz", "n_throttle_reasons"], "rows": per_rank},
            "gpu_launches": 6 * K,
            "gpu_launches_per_step": "corr: restage x2 levels + ONE two-level tcgen05 tile kernel; plan kernel (neighbors; its plan is reused by BA); window kernel (BA) + the one-thread kernel that mirrors BA's status word to mapped host memory -- profiles/r1_launches_stacked.csv",
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            try:
                out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                                     capture_output=True, text=True, timeout=900)
                ref = json.loads(out.stdout.strip().splitlines()[-1])
                line["cpu_baseline"] = ref["cpu_baseline"]
            except Exception as ex:  # the baseline is a reported number, not the product
                line["cpu_baseline"] = {"error": repr(ex)}
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    return run_reference(args) if args.impl == "reference" else run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
