#!/usr/bin/env python
"""bench.py -- one JSON line: frames/s of the reference's UNCHANGED tracker (DPVO.__call__, dpvo/dpvo.py:771-875) on this
repository's cuda_corr / cuda_ba / lietorch_backends, BASELINE.json configs[2].

A "step" is one frame through DPVO.__call__ with dpvo_configs/default.yaml (M = 384 patches/frame, fp16 maps) on a synthetic
512x384 5-FPS clip with depth prior and dynamic masks (harness/world.py: analytic ray casting, no graphics API).  The host code is
the reference's own, installed byte for byte under baseline/_ref by baseline/install_ref.py; only the three native modules it
imports by name are this repository's.  There are no pretrained weights on the box, so the network is harness.tracker.TeacherNet:
a seeded random-init VONet that runs its full patchify + update operator, with the flow head replaced by the synthetic world's
ground-truth flow + pixel noise (so the closed loop is meaningful and ATE can be reported).

  value     frames/s, frames already resident in HBM, CUDA events around exactly K frames after W warm-up frames (W >= 30: the
            tracker's own initialisation -- 10 frames + 12 updates -- must be over), max over ranks
  e2e       the same on a fresh tracker with every frame starting in pinned host memory (uploaded inside the timed region as
            dpvo_demo.py:55-60 does) and the newest pose read back after every frame
  ref_cuda  the same clip through the reference's OWN CUDA build for sm_100 (oracle/_ref/refgpu_*.so) in a subprocess
  hotpath   the replaced path alone at BASELINE configs[1] sizes (E = 45 312): DPVO.corr, neighbors, BA -- corr GB/s, BA iters/s
  roofline / roofline_ba / cpu_baseline / sharded   see DESIGN.md section 5

  python bench.py [--gpus N] [--steps K] [--warmup W]
  python bench.py --impl reference ...      the reference's own code for the replaced path on the host cores

Multi-GPU: a clip does not shard (replicas only); rank r tracks its own clip (world seed 2 + r), no data-path
collective, "scaling": "weak".  The path that does shard -- the loop-closure global BA -- is reported in "sharded".
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
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# algorithmic bytes per edge and level for fp16 maps (gmap patch + window footprint + coords +
# indices + output):  s*C*P^2 + s*C*F + 4*2*P^2 + 2*8 + s*49*P^2,  F0 = 100 px, F1 = 81 px
CORR_BYTES_L0 = 2 * 128 * 9 + 2 * 128 * 100 + 72 + 16 + 2 * 49 * 9   # 28 874
CORR_BYTES_L1 = 2 * 128 * 9 + 2 * 128 * 81 + 72 + 16 + 2 * 49 * 9    # 24 010
BA_BYTES_PER_EDGE_ITER = 116
# dram__bytes_read.sum + dram__bytes_write.sum of one DPVO.corr call at E = 45 312 (restage + corr_tc5_kernel<2>), ncu --set full
NCU_TRAFFIC_CORR = 630.4e6
# compulsory DRAM bytes of that call: both rings 120.3 MB + gmap 8.0 + output 80.0 + coords / indices 4.0
CORR_COMPULSORY_BYTES = 212.3e6
MIN_WARMUP = 30
WORKLOAD = ("configs[2]: unchanged DPVO.__call__ (dpvo/dpvo.py:771-875), dpvo_configs/default.yaml (M=384, REMOVAL_WINDOW 22, "
            "OPTIMIZATION_WINDOW 10, PATCH_LIFETIME 13, fp16 maps), synthetic 512x384 5-FPS clip with depth prior + dynamic masks "
            "(harness/world.py ray casting), TeacherNet = random-init VONet (full patchify + update operator) with ground-truth flow head")


def config(world):
    return {"workload": WORKLOAD, "parallelism": f"replicas x{world} (one clip per GPU, no collective)",
            "l2": "every frame streams new inputs: 590 KB image + 786 KB depth + the 150 MB per-level correlation volumes and the 113 MB "
                  "feature ring are larger than L2 (126 MB); no explicit flush"}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def make_hotpath_inputs(seed_offset=0, small=False, M=96):
    """numpy inputs of one hot-path step (seeded).  small=True: a reduced copy for the CPU tests."""
    syn = importlib.import_module(PKG + ".synthetic")
    if small:
        w = syn.make_corr_workload(seed=1 + 1000 * seed_offset, M=4, pmem=8, n_frames=10, ht=24, wd=32, lifetime=4, removal=6)
        s = syn.make_ba_scene(seed=2 + 1000 * seed_offset, n=10, M=4, edges="tracker", lifetime=4, removal=6)
        return w, s
    w = syn.make_corr_workload(seed=1 + 1000 * seed_offset, M=M)
    s = syn.make_ba_scene(seed=2 + 1000 * seed_offset, n=40, M=M, edges="tracker")
    assert w["E"] == len(s["kk"]) == 472 * M
    return w, s


make_inputs = make_hotpath_inputs   # name used by the CPU tests


class ClockSampler:
    """SM clock + clock-event (throttle) reasons sampled DURING the timed region: NVML in a thread, one sample per ms
    (the same counters `nvidia-smi --query-gpu=clocks.sm,clocks_event_reasons.*` prints, B200_PROFILING.md); `nvidia-smi -lms` is
    the fallback when NVML cannot be loaded."""
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
_REF_SHARED = None


def _ref_corr_worker(args):
    """one process: the reference's corr_forward (its own correlation_kernel.cu built as host code) on a slice of edges;
    returns its OWN compute time (module load and pool start-up are outside)"""
    lo, hi = args
    import torch
    torch.set_num_threads(1)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import build_ref
    rc = build_ref.load("ref_cuda_corr")
    if hi <= lo:
        return 0.0
    g = _REF_SHARED
    t = time.perf_counter()
    cs = [rc.forward(g["gmap"], fm, (g["coords"][:, lo:hi] / div).contiguous(), g["ii"][lo:hi], g["jj"][lo:hi], 3)[0].permute(0, 1, 3, 2, 4, 5)
          for fm, div in ((g["f0"], 1.0), (g["f1"], 4.0))]
    torch.stack(cs, -1).view(1, hi - lo, -1)   # dpvo/altcorr/correlation.py:30 + dpvo/dpvo.py:333
    return time.perf_counter() - t


class ReferenceHotPath:
    """the reference's own code for the replaced path of ONE frame on the host cores: corr on both levels (its CUDA kernel compiled
    as host code, oracle/_ref; edges split over `cores` processes) + neighbors + BA (2 iterations) on the full edge list.  The pool
    and the module loads happen once, outside every timed step; a step's corr time is the slowest worker's own compute time."""

    def __init__(self, M, cores):
        global _REF_SHARED
        import torch
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import build_ref
        self.torch, self.cores, self.M = torch, cores, M
        self.w, self.s = make_hotpath_inputs(0, M=M)
        self.E = self.w["E"]
        self.have_ref = all(os.path.exists(os.path.join(ROOT, "oracle", "_ref", n + ".so")) for n in ("ref_cuda_corr", "ref_cuda_ba"))
        w = self.w
        if self.have_ref:
            import multiprocessing as mp
            _REF_SHARED = dict(gmap=torch.from_numpy(w["gmap"]).half(), f0=torch.from_numpy(w["fmap0"]).half(),
                               f1=torch.from_numpy(w["fmap1"]).half(), coords=torch.from_numpy(w["coords"]),
                               ii=torch.from_numpy(w["ii1"]), jj=torch.from_numpy(w["jj1"]))
            self.pool = mp.get_context("fork").Pool(cores)
            self.pool.map(_ref_corr_worker, [(0, 0)] * cores)          # start the workers, load the module in each
            self.rb = build_ref.load("ref_cuda_ba")
            self.kind = "reference"
        else:
            import oracle as orc
            self.orc = orc
            self.kind = "port"

    def close(self):
        if self.have_ref:
            self.pool.close(); self.pool.join()

    def step(self, sample_edges):
        """-> (t_step, t_corr (scaled to E), t_ba, factor)"""
        torch, w, s, E = self.torch, self.w, self.s, self.E
        Es = min(E, sample_edges)
        T = lambda k: torch.from_numpy(s[k].copy())  # noqa: E731
        if self.have_ref:
            per = (Es + self.cores - 1) // self.cores
            slices = [(a, min(Es, a + per)) for a in range(0, Es, per)]
            t_corr = max(self.pool.map(_ref_corr_worker, slices, chunksize=1)) * (E / Es)
            torch.set_num_threads(self.cores)
            po, pa = T("poses"), T("patches")
            t = time.perf_counter()
            self.rb.neighbors(T("kk"), T("jj"))
            self.rb.forward(po, pa, T("intrinsics"), T("target"), T("weight"), T("lmbda"), T("ii"), T("jj"), T("kk"), s["t1"] - 10, s["t1"], 2)
            t_ba = time.perf_counter() - t
        else:
            orc = self.orc
            g = torch.from_numpy(w["gmap"]).half().float().numpy()
            t = time.perf_counter()
            for fm, div in ((w["fmap0"], 1.0), (w["fmap1"], 4.0)):
                f = torch.from_numpy(fm).half().float().numpy()
                orc.corr_forward(g, f, (w["coords"][:, :Es] / div).astype(np.float32), w["ii1"][:Es], w["jj1"][:Es], 3, mode=orc.F16EMU, nthreads=self.cores)
            t_corr = (time.perf_counter() - t) * (E / Es)
            t = time.perf_counter()
            orc.neighbors(s["kk"], s["jj"])
            orc.ba_forward(s["poses"], s["patches"], s["intrinsics"], s["target"], s["weight"], s["lmbda"], s["ii"], s["jj"], s["kk"],
                           s["t1"] - 10, s["t1"], 2, mode=orc.F32)
            t_ba = time.perf_counter() - t
        return t_corr + t_ba, t_corr, t_ba, E / Es


def time_python_ba(cores):
    """BASELINE configs[0]: the reference's own PyTorch BA (dpvo/ba.py::BA, imported unchanged from baseline/_ref over its own
    lietorch CPU ops built as host code, oracle/_ref/ref_lietorch.so) on a synthetic 10-frame window, 96 patches/frame, E = 9 600,
    2 Gauss-Newton iterations, on this machine's host cores.  -> dict or {"unavailable": why}"""
    try:
        import torch
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import build_ref
        lt = build_ref.load("ref_lietorch")
        if lt is None or not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "dpvo")):
            return {"unavailable": "oracle/_ref/ref_lietorch.so or baseline/_ref missing"}
        sys.modules["lietorch_backends"] = lt
        sys.modules["cuda_ba"] = build_ref.load("ref_cuda_ba")
        sys.modules["cuda_corr"] = build_ref.load("ref_cuda_corr")
        shim = os.path.join(ROOT, "harness", "shims")
        if shim not in sys.path:
            sys.path.append(shim)
        sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
        pyba = importlib.import_module("dpvo.ba")
        lie = importlib.import_module("dpvo.lietorch")
        syn = importlib.import_module(PKG + ".synthetic")
        torch.set_num_threads(cores)
        T = torch.from_numpy
        s = syn.make_ba_scene(seed=0, n=10, M=96, edges="all")
        cx, cy = s["intrinsics"][0, 2], s["intrinsics"][0, 3]
        bounds = [-64, -64, 2 * cx + 64, 2 * cy + 64]
        ts = []
        for rep in range(6):
            G, pat = lie.SE3(T(s["poses"].copy())[None]), T(s["patches"].copy())[None]
            t = time.perf_counter()
            for _ in range(2):
                G, pat = pyba.BA(G, pat, T(s["intrinsics"])[None], T(s["target"])[None], T(s["weight"])[None], 1e-4, T(s["ii"]), T(s["jj"]),
                                 T(s["kk"]), bounds, ep=1.0, fixedp=s["t0"], patches_est=T(s["patches_est"])[None])
            ts.append(time.perf_counter() - t)
        ms = 1e3 * float(np.median(ts[2:]))
        return {"ms_per_2_iterations": ms, "iters_per_s": 2e3 / ms, "cores": cores, "E": int(len(s["kk"])),
                "what": "dpvo/ba.py::BA (reference PyTorch CPU path, BASELINE configs[0]) over the reference's lietorch_cpu.cpp"}
    except Exception as ex:
        return {"unavailable": repr(ex)[:300]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    ref = ReferenceHotPath(args.hotpath_m, cores)
    n_steps = args.warmup + args.steps
    # bounded sample: the whole run should end within a few minutes -> at most ~150 s of corr over all steps
    t_probe, t_corr_probe, _, _ = ref.step(max(cores * 16, 2048))
    per_edge = t_corr_probe / ref.E
    budget_edges = int(150.0 / max(n_steps, 1) / max(per_edge, 1e-9))
    sample_edges = ref.E if budget_edges >= ref.E else max(cores * 16, min(budget_edges, ref.E))
    times = []
    for it in range(n_steps):
        r = ref.step(sample_edges)
        if it >= args.warmup:
            times.append(r)
    ref.close()
    t_step = float(np.mean([t[0] for t in times]))
    factor = float(times[0][3])
    sample = (f"hot path of ONE frame at default.yaml size (E = {ref.E}): corr on {min(sample_edges, ref.E)} of {ref.E} edges x 2 levels over "
              f"{cores} processes (slowest worker's own compute time, x{factor:.2f}); neighbors + BA(2 iters) on the full edge list; "
              + ("reference's own correlation_kernel.cu / ba_cuda.cu / ba.cpp compiled as host code (oracle/_ref)" if ref.have_ref
                 else "oracle port (oracle/liboracle.so), fp16-emulating corr")
              + "; the network and the Python host code of a frame are NOT included (the reference has no CPU build of the tracker: dpvo/dpvo.py:20)")
    line = {
        "impl": "reference", "metric": "dpvo_frames_per_s", "value": 1.0 / t_step, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f16 maps / f32 accumulate+geometry", "data": "synthetic",
        "config": config(args.gpus), "extrapolated": factor > 1.0, "factor": factor,
        "n_gpus_note": "whole-host aggregate: the same host cores would serve all N clips, so the value does not grow with --gpus",
        "cpu_baseline": {"value": 1.0 / t_step, "unit": "frames/s", "cores": cores, "kind": ref.kind, "sample": sample,
                         "corr_s_per_step": float(np.mean([t[1] for t in times])), "ba_s_per_step": float(np.mean([t[2] for t in times]))},
        "e2e": {"value": 1.0 / t_step, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if args.python_ba:
        line["cpu_baseline"]["python_ba_config0"] = time_python_ba(cores)
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------------------
def hotpath_microbench(pkg, dev, rank, steps, M=96):
    """the replaced path alone (BASELINE configs[1] sizes at M = 96): DPVO.corr as one stacked call and as the drop-in two-call
    form, neighbors, BA (10 free poses, 2 GN iterations); 512 MB fill between steps (outside the event intervals)"""
    import torch
    cc, cb = pkg.native_modules()
    altcorr = importlib.import_module(PKG + ".altcorr.correlation")
    lib = pkg.capi()
    w, s = make_hotpath_inputs(rank, M=M)
    E = w["E"]
    to = lambda a, dt=None: (torch.from_numpy(np.ascontiguousarray(a)).to(dt) if dt else torch.from_numpy(np.ascontiguousarray(a))).to(dev)  # noqa: E731
    D = dict(gmap=to(w["gmap"], torch.half), f0=to(w["fmap0"], torch.half), f1=to(w["fmap1"], torch.half), c0=to(w["coords"]),
             ii1=to(w["ii1"]), jj1=to(w["jj1"]), poses=to(s["poses"]), patches=to(s["patches"]), intr=to(s["intrinsics"]),
             target=to(s["target"]), weight=to(s["weight"]), lmbda=to(s["lmbda"]), ii=to(s["ii"]), jj=to(s["jj"]), kk=to(s["kk"]))
    t0, t1 = int(s["t1"]) - 10, int(s["t1"])
    poses, patches = D["poses"].clone(), D["patches"].clone()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    seg = []

    def step(timed):
        poses.copy_(D["poses"]); patches.copy_(D["patches"])
        flush.fill_(1)
        e = [ev() for _ in range(6)]
        e[0].record()
        corr = altcorr.corr_stacked(D["gmap"], (D["f0"], D["f1"]), D["c0"], D["ii1"], D["jj1"], 3)
        e[1].record()
        ix, jx = cb.neighbors(D["kk"], D["jj"])
        e[2].record()
        cb.forward(poses, patches, D["intr"], D["target"], D["weight"], D["lmbda"], D["ii"], D["jj"], D["kk"], t0, t1, 2)
        e[3].record()
        flush.fill_(1)
        e[4].record()   # the drop-in two-call form, exactly dpvo/dpvo.py:331-333
        c1 = altcorr.corr(D["gmap"], D["f0"], D["c0"] / 1, D["ii1"], D["jj1"], 3)
        c2 = altcorr.corr(D["gmap"], D["f1"], D["c0"] / 4, D["ii1"], D["jj1"], 3)
        two = torch.stack([c1, c2], -1).view(1, E, -1)
        e[5].record()
        if timed:
            seg.append(e)
        return corr, two

    for _ in range(3):
        corr, two = step(False)
    torch.cuda.synchronize()
    assert torch.equal(corr, two), "stacked corr != two-call form"
    l0 = lib.wv_launch_count()
    for _ in range(steps):
        step(True)
    torch.cuda.synchronize()
    launches = (lib.wv_launch_count() - l0) / steps
    t = np.array([[e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), e[2].elapsed_time(e[3]), e[4].elapsed_time(e[5])] for e in seg]).mean(0)
    # grid-wide barrier latency of a BA-shaped cooperative grid (for roofline_ba)
    lib.wv_gridsync_probe.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for iters in (10, 1010):
        lib.wv_gridsync_probe(0, iters, st)
    ea, eb, ec = ev(), ev(), ev()
    ea.record(); lib.wv_gridsync_probe(0, 10, st); eb.record(); lib.wv_gridsync_probe(0, 1010, st); ec.record()
    torch.cuda.synchronize()
    t_sync_us = 1e3 * (eb.elapsed_time(ec) - ea.elapsed_time(eb)) / 1000.0
    del flush
    return dict(E=E, M=M, corr_stacked_ms=float(t[0]), neighbors_ms=float(t[1]), ba_2iters_ms=float(t[2]), corr_two_call_ms=float(t[3]),
                launches_per_step=launches, t_grid_sync_us=float(t_sync_us))


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
    lib = pkg.capi()
    lib.wv_launch_count.restype = ctypes.c_ulonglong
    assert lib.wv_device_supported() == 1, "libwvb200 has sm_100a code only"
    from harness import tracker as H
    from harness.world import World
    H.load_tracker("b200")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([float(v)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    W, K, B = max(args.warmup, MIN_WARMUP), args.steps, args.breakdown_frames
    wld = World(seed=2 + rank, n_frames=W + K + B, device=dev)
    frames = [wld.frame(t) for t in range(W + K + B)]
    torch.cuda.synchronize()

    # ---- headline: frames resident in HBM ------------------------------------------------------------------------------------
    class TimedRegionProbe:   # started by the runner at the first timed frame, stopped right after the last
        def __init__(self):
            self.cs, self.l0, self.l1, self.clocks = ClockSampler(local), 0, 0, None

        def start(self):
            self.l0 = lib.wv_launch_count()
            self.cs.start()

        def stop(self):
            self.clocks = self.cs.stop()
            self.l1 = lib.wv_launch_count()
    probe = TimedRegionProbe()
    r = H.run_clip_windows(wld, "b200", W, K, B, frames=frames, barrier=barrier, sampler=probe)
    clocks = probe.clocks
    t_max = max_over_ranks(r["timed_ms"])
    launches = probe.l1 - probe.l0          # kernels libwvb200.so launched inside the timed region (every launch site counts itself)
    launches_per_frame = launches / float(K)

    # ---- e2e: frames start in pinned host memory, pose read back per frame ------------------------------------------------------
    r2 = H.run_clip_windows(wld, "b200", W, K, 0, frames=frames, frames_on_host=True, barrier=barrier, net=r["net"])
    te_max = max_over_ranks(r2["timed_ms"])

    # ---- optional integration: the update operator's SoftAgg modules replaced by this repository's fused
    #      blocks.SoftAgg.  NOT the headline: the headline runs the reference's network untouched. ------------------
    fused = None
    if world == 1 and not args.quick:
        try:
            Kf = min(K, 100)
            r3 = H.run_clip_windows(wld, "b200", W, Kf, 0, frames=frames, barrier=barrier, fused_softagg=True)
            r4 = H.run_clip_windows(wld, "b200", W, Kf, 0, frames=frames, barrier=barrier, fused_softagg="all")
            fused = {"value": Kf / (r3["timed_ms"] * 1e-3), "unit": "frames/s", "steps": Kf, "ms_per_frame": r3["timed_ms"] / Kf,
                     "all_three_switches": {"value": Kf / (r4["timed_ms"] * 1e-3), "ms_per_frame": r4["timed_ms"] / Kf, "ate_rmse_m": r4["ate"],
                                            "what": "+ DPVO.corr as ONE altcorr.corr_stacked call (dpvo/dpvo.py:325-333) and dpvo.dpvo's `pops` "
                                                    "module replaced by this repository's fused projective_ops (transform / flow_mag / point_cloud)"},
                     "ate_rmse_m": r3["ate"], "keyframes": r3["keyframes"],
                     "what": "same clip and loop; dpvo.net's two SoftAgg modules are wild-video-3d-reconstruction_b200.blocks.SoftAgg (group "
                             "plan + one fused scatter-softmax / scatter-sum kernel, same weights) instead of torch.unique + the "
                             "torch_scatter stand-in; optional one-line switch, not part of `value`"}
        except Exception as ex:
            fused = {"error": repr(ex)[:300]}

    # ---- the replaced path alone (sub-metrics + rooflines) -----------------------------------------------------------------------
    hp = hotpath_microbench(pkg, dev, rank, args.hotpath_steps, M=96)
    hp384 = hotpath_microbench(pkg, dev, rank, max(3, args.hotpath_steps // 2), M=384) if rank == 0 and not args.quick else None

    per_rank = None
    if world > 1:
        mine = torch.tensor([r["timed_ms"] / K, r2["timed_ms"] / K, clocks["sm_mhz"] or 0.0, float(len(clocks["reasons"]))], device=dev, dtype=torch.float64)
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = [[round(float(v), 4) for v in x.tolist()] for x in allr]
        clocks["sm_mhz_min_over_ranks"] = min(x[2] for x in per_rank)

    sharded = None
    if args.sharded:
        try:
            from tools import global_ba as gba
            sharded = gba.bench_sharded(dev, rank, world, quick=args.quick)
        except Exception as ex:
            sharded = {"error": repr(ex)[:300]}

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        corr_ms = hp["corr_stacked_ms"]
        E = hp["E"]
        alg = (CORR_BYTES_L0 + CORR_BYTES_L1) * E
        achieved = alg / (corr_ms * 1e-3) / 1e9
        ba_ms = hp["ba_2iters_ms"]
        n_sync = 3 * 2                        # 3 grid-wide barriers per Gauss-Newton iteration (csrc/ba_window.cu), 2 iterations
        t_model_us = n_sync * hp["t_grid_sync_us"] + 2 * 6.0   # + the 60x60 in-smem Cholesky + substitutions' dependent chain, ~6 us / iteration
        line = {
            "metric": "dpvo_frames_per_s", "value": world * K / (t_max * 1e-3), "unit": "frames/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": t_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16 maps / f32 accumulate+geometry", "data": "synthetic", "config": config(world),
            "tracker": {"keyframes": r["keyframes"], "edges_at_end": r["edges"], "ate_rmse_m": r["ate"], "ate_rmse_m_e2e_run": r2["ate"],
                        "frames_total": r["frames_total"], "backend_modules": [os.path.basename(m.__file__) for m in H._loaded["native"]]},
            "ms_breakdown": r["ms_breakdown"],
            "ms_breakdown_note": f"{B} frames after the timed region with a device synchronisation around every phase (attribution: the "
                                 "phases sum to more than an un-instrumented frame); ba / corr / reproject+point_cloud (lietorch) are this repository's kernels",
            "e2e": {"value": world * K / (te_max * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": int(r2["h2d_bytes_per_frame"]),
                    "d2h_bytes_per_step": int(r2["d2h_bytes_per_frame"]),
                    "api": "DPVO.__call__ per frame: image/depth/mask/intrinsics uploaded from pinned host memory (dpvo_demo.py:55-60), newest pose read back"},
            "hotpath": {"workload": f"BASELINE configs[1]: M=96, pmem=36, E={E}, radius 3, fp16 maps; BA 10 free poses, 2 GN iterations",
                        **{k: hp[k] for k in ("corr_stacked_ms", "corr_two_call_ms", "neighbors_ms", "ba_2iters_ms", "launches_per_step")},
                        "frames_per_s": 1e3 / (hp["corr_stacked_ms"] + hp["neighbors_ms"] + hp["ba_2iters_ms"]),
                        "dropin_frames_per_s": 1e3 / (hp["corr_two_call_ms"] + hp["neighbors_ms"] + hp["ba_2iters_ms"])},
            "hotpath_m384": ({k: hp384[k] for k in ("E", "corr_stacked_ms", "corr_two_call_ms", "neighbors_ms", "ba_2iters_ms")} if hp384 else None),
            "corr_gbs": achieved,
            "ba_iters_per_s": 2.0 / (ba_ms * 1e-3),
            "roofline": {"kernel": "DPVO.corr at configs[1]: both levels, stacked (restage_kernel x2 + corr_tc5_kernel<2>, fp16 maps)",
                         "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC_CORR, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg,
                         "traffic_source": "dram__bytes_read+write of restage_kernel x2 + corr_tc5_kernel<2>, ncu --set full, profiles/r2_corr_stacked_ncu_summary.txt",
                         "dram_gbs_from_traffic": NCU_TRAFFIC_CORR / (corr_ms * 1e-3) / 1e9,
                         "dram_frac_of_peak": NCU_TRAFFIC_CORR / (corr_ms * 1e-3) / 1e9 / peak,
                         "compulsory_bytes": CORR_COMPULSORY_BYTES,
                         "compulsory_frac_of_peak": CORR_COMPULSORY_BYTES / (corr_ms * 1e-3) / 1e9 / peak,
                         "l2_to_sm_tile_gbs": 2 * E * 28.0e3 / (corr_ms * 1e-3) / 1e9,
                         "note": "frac > 1 is NOT an HBM-efficiency claim: the SURVEY 8(d) per-edge bytes count every edge's window as an HBM "
                                 "read, neighbouring windows overlap ~10x and the L2 serves them.  Honest companions: dram_frac_of_peak "
                                 "(measured DRAM traffic) and compulsory_frac_of_peak; the tile kernel's own ceiling is the L2->SM tile stream"},
            "roofline_ba": {"kernel": "ba_window_kernel: fastba.BA, 10 free poses, 2 GN iterations, E = %d" % E, "bound": "latency",
                            "model": "t_min = n_grid_syncs * t_sync + dependent Cholesky chain (SURVEY 8d: max(bytes/HBM, flops/FP32, n_dependent * t))",
                            "n_grid_syncs": n_sync, "t_grid_sync_us": hp["t_grid_sync_us"], "t_model_us": t_model_us,
                            "achieved_us": 1e3 * ba_ms, "frac": t_model_us / (1e3 * ba_ms), "unit": "us",
                            "hbm_bytes": BA_BYTES_PER_EDGE_ITER * E * 2, "hbm_frac_of_peak": BA_BYTES_PER_EDGE_ITER * E * 2 / (ba_ms * 1e-3) / 1e9 / peak},
            "per_rank": {"columns": ["ms_per_frame", "e2e_ms_per_frame", "sm_mhz", "n_throttle_reasons"], "rows": per_rank},
            "gpu_launches": int(launches),
            "gpu_launches_per_frame": launches_per_frame,
            "gpu_launches_note": "kernels launched by libwvb200.so (every launch site counts itself, wv_launch_count): per frame 2 x (restage + "
                                 "tcgen05 tile kernel) for corr, the plan kernel (neighbors), the BA window kernel + status mirror, ~12 lietorch "
                                 "group ops per reprojection / point cloud, 4 patchify gathers",
            "sharded": sharded,
            "with_fused_softagg": fused,
            "clocks": clocks,
        }
        if world == 1 and not args.no_ref_cuda:
            line["ref_cuda"] = ref_cuda_subprocess(W, K)
            if line["ref_cuda"].get("value"):
                line["ref_cuda"]["speedup"] = line["value"] / line["ref_cuda"]["value"]
        elif world > 1:
            line["ref_cuda"] = {"skipped": "REF-CUDA is run beside the N=1 line only"}
        if world == 1 and not args.no_cpu:
            try:
                out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "1", "--warmup", "0", "--python-ba"],
                                     capture_output=True, text=True, timeout=900)
                ref = json.loads(out.stdout.strip().splitlines()[-1])
                line["cpu_baseline"] = ref["cpu_baseline"]
                line["cpu_baseline"]["extrapolated"], line["cpu_baseline"]["factor"] = ref["extrapolated"], ref["factor"]
            except Exception as ex:  # the baseline is a reported number, not the product
                line["cpu_baseline"] = {"error": repr(ex)[:300]}
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def ref_cuda_subprocess(W, K):
    """the same clip through the reference's own CUDA build (one backend per process)"""
    need = [os.path.join(ROOT, "oracle", "_ref", f"refgpu_{n}.so") for n in ("cuda_corr", "cuda_ba", "lietorch_backends")]
    if not all(os.path.exists(p) for p in need):
        return {"skipped": "oracle/_ref/refgpu_*.so absent (built in the authoring container by oracle/build_ref_gpu.py)"}
    try:
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "refcuda", "--steps", str(K), "--warmup", str(W)],
                             capture_output=True, text=True, timeout=1200)
        return json.loads(out.stdout.strip().splitlines()[-1])
    except Exception as ex:
        return {"error": repr(ex)[:300]}


def run_refcuda(args):
    """internal leg: the tracker over REF-CUDA, same clip / windows as the headline; prints a small JSON object"""
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import torch
    from harness import tracker as H
    from harness.world import World
    H.load_tracker("refcuda")
    W, K = max(args.warmup, MIN_WARMUP), args.steps
    wld = World(seed=2, n_frames=W + K, device="cuda")
    r = H.run_clip_windows(wld, "refcuda", W, K, 0)
    out = {"value": K / (r["timed_ms"] * 1e-3), "unit": "frames/s", "ms_per_frame": r["timed_ms"] / K, "ate_rmse_m": r["ate"],
           "keyframes": r["keyframes"], "edges_at_end": r["edges"],
           "what": "same clip, same host code, the reference's own cuda_corr / cuda_ba / lietorch_backends compiled unedited for sm_100 (oracle/build_ref_gpu.py)"}
    real_stdout.write(json.dumps(out) + "\n")
    real_stdout.flush()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=40)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "refcuda"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip the REF-CUDA tracker leg")
    ap.add_argument("--no-sharded", dest="sharded", action="store_false", help="skip the sharded global-BA leg")
    ap.add_argument("--quick", action="store_true", help="smaller sub-benchmarks (development)")
    ap.add_argument("--python-ba", action="store_true", help="reference arm: also time dpvo/ba.py (configs[0])")
    ap.add_argument("--breakdown-frames", type=int, default=40)
    ap.add_argument("--hotpath-steps", type=int, default=10)
    ap.add_argument("--hotpath-m", type=int, default=384, help="reference arm: patches per frame of the hot-path workload")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.impl == "refcuda":
        return run_refcuda(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
