#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY -- builds the reference's own sources as CPU code.

Compiles, from where they lie under /root/reference (never copied into the repo):

  ref_cuda_corr      dpvo/altcorr/correlation.cpp + correlation_kernel.cu
  ref_cuda_ba        dpvo/fastba/ba.cpp + ba_cuda.cu            (Eigen: thirdparty/eigen)
  ref_lietorch       dpvo/lietorch/src/lietorch.cpp + lietorch_cpu.cpp

into oracle/_ref/*.so (git-ignored).  The .cu files are compiled by g++ as host code:
oracle/ref_shim/cuda_host_emul.h maps the CUDA keywords they use onto serial host
execution, and this script performs the one edit a macro cannot express, rewriting

    kernel<<<grid, block>>>(args...)   ->   HOST_LAUNCH(grid, block, kernel, args...)

in a temporary copy that is deleted after compilation.  No arithmetic line is touched,
so the resulting modules ARE the reference's fp32/fp16 arithmetic (atomics applied in
thread order), and are used to (1) validate the C restatement in oracle/*.c and
(2) generate the golden vectors under tests/golden/ (oracle/make_golden.py).

The reference's own build system (setup.py) is not run.  /root/reference is absent on
the GPU box: only the prebuilt oracle/_ref/*.so travel there and nothing needs them at
run time except optional cross-checks that skip when they are missing.
"""
import os
import re
import shutil
import subprocess
import sys
import sysconfig
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("WV_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")
SHIM = os.path.join(HERE, "ref_shim")

_LAUNCH = re.compile(r"([A-Za-z_]\w*(?:<[^<>;(){}]*>)?)\s*<<<\s*(.*?)\s*,\s*([A-Za-z_0-9()* +]+?)\s*>>>\s*\(")


def rewrite_launches(src: str) -> str:
    """kernel<<<G, B>>>(  ->  HOST_LAUNCH(G, B, kernel,   (argument list is left as is)."""
    return _LAUNCH.sub(lambda m: f"HOST_LAUNCH({m.group(2)}, {m.group(3)}, {m.group(1)}, ", src)


def _torch_flags():
    import torch
    from torch.utils import cpp_extension as ce
    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}"]
    libdir = ce.library_paths()[0]
    link = [f"-L{libdir}", f"-Wl,-rpath,{libdir}", "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python"]
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    return inc, link, [f"-D_GLIBCXX_USE_CXX11_ABI={abi}"]


def _compile(name, wrapper_srcs, extra_inc, tmp):
    """one translation unit per reference source file (as in the reference's own build)."""
    inc, link, defs = _torch_flags()
    out = os.path.join(OUT, f"{name}.so")
    objs = []
    for i, src in enumerate(wrapper_srcs):
        wrapper = os.path.join(tmp, f"{name}_tu{i}.cpp")
        with open(wrapper, "w") as f:
            f.write(src)
        obj = wrapper[:-4] + ".o"
        cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-c", "-w", "-fopenmp",
               f"-DTORCH_EXTENSION_NAME={name}", "-DTORCH_API_INCLUDE_EXTENSION_H",
               f"-I{SHIM}", f"-I{SHIM}/cuda_stub"] + [f"-I{p}" for p in extra_inc] + inc + defs + \
              [wrapper, "-o", obj]
        print("[build_ref]", name, f"tu{i}", flush=True)
        subprocess.check_call(cmd)
        objs.append(obj)
    subprocess.check_call(["g++", "-shared", "-fopenmp"] + objs + ["-o", out] + link)
    return out


def build(force=False):
    if not os.path.isdir(REF):
        print(f"[build_ref] {REF} not present; keeping prebuilt oracle/_ref as is")
        return False
    os.makedirs(OUT, exist_ok=True)
    want = ["ref_cuda_corr", "ref_cuda_ba", "ref_lietorch"]
    if not force and all(os.path.exists(os.path.join(OUT, f"{w}.so")) for w in want):
        return True
    tmp = tempfile.mkdtemp(prefix="wv_ref_build_")
    try:
        # --- cuda_corr ---------------------------------------------------------------
        ker = rewrite_launches(open(f"{REF}/dpvo/altcorr/correlation_kernel.cu").read())
        open(os.path.join(tmp, "corr_kernel_host.inc"), "w").write(ker)
        _compile("ref_cuda_corr", [f'''
#include "cuda_host_emul.h"
#include "{tmp}/corr_kernel_host.inc"
''', f'''
#include "cuda_host_emul.h"
#include "{REF}/dpvo/altcorr/correlation.cpp"
'''], [], tmp)
        # --- cuda_ba -----------------------------------------------------------------
        ker = rewrite_launches(open(f"{REF}/dpvo/fastba/ba_cuda.cu").read())
        open(os.path.join(tmp, "ba_kernel_host.inc"), "w").write(ker)
        _compile("ref_cuda_ba", [f'''
#include "cuda_host_emul.h"
#include "{tmp}/ba_kernel_host.inc"
''', f'''
#include "cuda_host_emul.h"
#include "{REF}/dpvo/fastba/ba.cpp"
'''], [f"{REF}/thirdparty/eigen"], tmp)
        # --- lietorch (CPU half only; the *_gpu symbols stay unresolved and are never
        #     reached for CPU tensors -- the module is imported with RTLD_LAZY) -------
        _compile("ref_lietorch", [f'''
#include "cuda_host_emul.h"
#undef kCUDA
#include "{REF}/dpvo/lietorch/src/lietorch_cpu.cpp"
''', f'''
#include "cuda_host_emul.h"
#undef kCUDA
#include "{REF}/dpvo/lietorch/src/lietorch.cpp"
'''], [f"{REF}/dpvo/lietorch/include", f"{REF}/thirdparty/eigen"], tmp)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return True


def load(name):
    """import oracle/_ref/<name>.so as a module (None when it was never built)."""
    import importlib.util
    import torch  # noqa: F401  (libtorch must be loaded first)
    path = os.path.join(OUT, f"{name}.so")
    if not os.path.exists(path):
        return None
    old = sys.getdlopenflags()
    sys.setdlopenflags(os.RTLD_LAZY | os.RTLD_LOCAL)
    try:
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.setdlopenflags(old)
    return mod


if __name__ == "__main__":
    build(force="--force" in sys.argv)
