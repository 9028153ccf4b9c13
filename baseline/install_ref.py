#!/usr/bin/env python
"""MEASUREMENT / TEST INFRASTRUCTURE ONLY -- installs the UNMODIFIED reference under baseline/_ref/ (git-ignored, travels to the
GPU box with the gpurun snapshot; /root/reference itself does not exist there).

  baseline/_ref/dpvo/...            the reference's Python host code (dpvo/*.py, dpvo/{altcorr,fastba,lietorch,loop_closure}/*.py),
                                    copied byte for byte -- the tracker this repository's kernels are a drop-in for
  baseline/_ref/dpvo_configs/*.yaml its tracker presets (default.yaml, dpvo_2k.yaml, fast.yaml, tum_default.yaml)
  oracle/_ref/ref_*.so              the reference's .cu/.cpp compiled as host code      (oracle/build_ref.py)
  oracle/_ref/refgpu_*.so           the reference's .cu/.cpp compiled by nvcc for sm_100 (oracle/build_ref_gpu.py): "REF-CUDA",
                                    cuda_corr + cuda_ba + lietorch_backends

The reference's own build system (setup.py) is not run: it points at a non-existent thirdparty/eigen-3.4.0 and does not compile
against torch 2.11 without the two source-compat shims of oracle/ref_shim/cuda_gpu_compat.h.
Nothing is copied into tracked paths; nothing in the product imports baseline/_ref.

    python baseline/install_ref.py [--no-gpu-build] [--force]
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("WV_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")


def copy_host_code():
    n = 0
    for top in ("dpvo", "dpvo_configs"):
        for d, _, files in os.walk(os.path.join(REF, top)):
            for f in files:
                if not f.endswith((".py", ".yaml")):
                    continue
                src = os.path.join(d, f)
                dst = os.path.join(DST, os.path.relpath(src, REF))
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                if not (os.path.exists(dst) and filecmp.cmp(src, dst, shallow=False)):
                    shutil.copyfile(src, dst)
                n += 1
    return n


def install(gpu_build=True, force=False):
    if not os.path.isdir(REF):
        print(f"[install_ref] {REF} not present; keeping baseline/_ref as is")
        return False
    n = copy_host_code()
    print(f"[install_ref] {n} reference host files under {DST}")
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import build_ref
    build_ref.build(force)
    if gpu_build:
        import build_ref_gpu
        build_ref_gpu.build(force)
    return True


if __name__ == "__main__":
    install("--no-gpu-build" not in sys.argv, "--force" in sys.argv)
