"""Builds the native pieces IN-TREE (the .so files travel to the GPU box with the snapshot):

  libwvb200.so        nvcc, sm_100a only, C ABI of include/wvb200.h, no torch
  cuda_corr.*.so      g++, torch extension with the reference's module name / m.def table (correlation.cpp:57-63)
  cuda_ba.*.so        g++, torch extension with the reference's module name / m.def table (ba.cpp:236-241)
  cuda_agg.*.so       g++, torch extension for the update operator's SoftAgg (the reference uses torch_scatter there, blocks.py:41-43)
  lietorch_backends.*.so   g++, torch extension with the reference's module name / m.def table (lietorch.cpp:286-316)

python build.py [--force] [--verbose]
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
NVCC = os.path.join(CUDA_HOME, "bin", "nvcc")
HOST_CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"

CU_SOURCES = ["abi.cu", "host_entry.cu", "corr_generic.cu", "corr_sm100.cu", "ba_plan.cu", "ba_window.cu", "ba_global.cu", "ba_band.cu", "pops.cu", "agg.cu", "colmap_text.cu", "lie_ops.cu", "neighbors_static.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
              "-ccbin", HOST_CXX, "--expt-relaxed-constexpr"]
EXT_SUFFIX = sysconfig.get_config_var("EXT_SUFFIX")


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def build_lib(force=False, verbose=False):
    out = os.path.join(HERE, "libwvb200.so")
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))] + \
              [os.path.join(ROOT, "include", "wvb200.h")]
    objs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(HERE, "build", src.replace(".cu", ".o"))
        if force or _newer(o, [s] + headers):
            _run([NVCC] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o], verbose)
        objs.append(o)
    if force or _newer(out, objs):
        _run([NVCC, "-shared", "-o", out] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", HOST_CXX,
                                                    "-lcudart_static", "-ldl", "-lrt", "-lpthread"], verbose)
    return out


def build_ext(name, force=False, verbose=False):
    import torch
    from torch.utils import cpp_extension as ce
    src = os.path.join(CSRC, f"bind_{name}.cpp")
    out = os.path.join(HERE, name + EXT_SUFFIX)
    deps = [src, os.path.join(ROOT, "include", "wvb200.h"), os.path.join(CSRC, "bind_common.h")]
    if not (force or _newer(out, deps)):
        return out
    libdir = ce.library_paths()[0]
    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", f"-I{CUDA_HOME}/include",
                                                    f"-I{os.path.join(ROOT, 'include')}"]
    cmd = [HOST_CXX, "-O2", "-std=c++17", "-fPIC", "-shared", "-w", f"-DTORCH_EXTENSION_NAME={name}",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"] + inc + \
          [src, "-o", out, f"-L{HERE}", "-lwvb200", "-Wl,-rpath,$ORIGIN", f"-L{libdir}", f"-Wl,-rpath,{libdir}",
           "-ltorch", "-ltorch_cpu", "-lc10", "-lc10_cuda", "-ltorch_python"]
    _run(cmd, verbose)
    return out


def build_all(force=False, verbose=False):
    lib = build_lib(force, verbose)
    exts = [build_ext(n, force, verbose) for n in ("cuda_corr", "cuda_ba", "cuda_agg", "lietorch_backends")]
    return [lib] + exts


if __name__ == "__main__":
    for p in build_all("--force" in sys.argv, "--verbose" in sys.argv):
        print(p)
