#!/usr/bin/env python
"""MEASUREMENT INFRASTRUCTURE ONLY -- the reference's own CUDA build for sm_100 ("REF-CUDA").

Compiles, from where they lie under /root/reference (never copied into the repo), with nvcc for sm_100:

  refgpu_cuda_corr   dpvo/altcorr/correlation.cpp + correlation_kernel.cu
  refgpu_cuda_ba     dpvo/fastba/ba.cpp + ba_cuda.cu            (Eigen: thirdparty/eigen)
  refgpu_lietorch_backends   dpvo/lietorch/src/{lietorch.cpp, lietorch_cpu.cpp, lietorch_gpu.cu}   (Eigen: thirdparty/eigen)

into oracle/_ref/*.so (git-ignored; they travel to the GPU box with the snapshot).  The only additions are the two
source-compatibility shims of oracle/ref_shim/cuda_gpu_compat.h (pre-included); no line of the reference is edited.  The
reference's own build system (setup.py) is not run.  Used by tools/ref_cuda_compare.py: the reference's kernels and this
repository's, same inputs, same B200, same harness -- and a full-size parity check of this build against the reference's own
GPU results.  Nothing in the product, the tests' pass/fail or bench.py depends on these modules.

    python oracle/build_ref_gpu.py [--force]
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("WV_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")
SHIM = os.path.join(HERE, "ref_shim")
NVCC = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "nvcc")


def _flags(name):
    import torch
    from torch.utils import cpp_extension as ce
    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", "-I/usr/local/cuda/include"]
    libdir = ce.library_paths()[0]
    link = [f"-L{libdir}", f"-Xlinker", f"-rpath={libdir}", "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10", "-lc10_cuda", "-ltorch_python"]
    common = ["-gencode", "arch=compute_100,code=sm_100", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-w", "--expt-relaxed-constexpr",
              "--expt-extended-lambda", f"-DTORCH_EXTENSION_NAME={name}", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-include", os.path.join(SHIM, "cuda_gpu_compat.h")]
    return inc, link, common


def _build(name, sources, extra_inc, force):
    out = os.path.join(OUT, f"{name}.so")
    if os.path.exists(out) and not force:
        return out
    inc, link, common = _flags(name)
    objs = []
    for i, src in enumerate(sources):
        obj = os.path.join(OUT, f".{name}_tu{i}.o")
        print("[build_ref_gpu]", name, os.path.basename(src), flush=True)
        subprocess.check_call([NVCC] + common + ["-x", "cu", "-c", src, "-o", obj] + inc + [f"-I{p}" for p in extra_inc])
        objs.append(obj)
    subprocess.check_call([NVCC, "-shared", "-o", out] + objs + link)
    for o in objs:
        os.remove(o)
    return out


def build(force=False):
    if not os.path.isdir(REF):
        print(f"[build_ref_gpu] {REF} not present; keeping prebuilt oracle/_ref as is")
        return False
    os.makedirs(OUT, exist_ok=True)
    _build("refgpu_cuda_corr", [f"{REF}/dpvo/altcorr/correlation_kernel.cu", f"{REF}/dpvo/altcorr/correlation.cpp"], [], force)
    _build("refgpu_cuda_ba", [f"{REF}/dpvo/fastba/ba_cuda.cu", f"{REF}/dpvo/fastba/ba.cpp"], [f"{REF}/thirdparty/eigen"], force)
    # lietorch_backends (dpvo/lietorch/group_ops.py:1 imports it by name; the tracker's reprojection runs on it)
    lt = f"{REF}/dpvo/lietorch"
    _build("refgpu_lietorch_backends", [f"{lt}/src/lietorch_gpu.cu", f"{lt}/src/lietorch_cpu.cpp", f"{lt}/src/lietorch.cpp"],
           [f"{lt}/include", f"{REF}/thirdparty/eigen"], force)
    return True


if __name__ == "__main__":
    build(force="--force" in sys.argv)
