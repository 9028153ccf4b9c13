"""loads oracle/_ref/refgpu_<name>.so (the reference's own sources compiled by nvcc for sm_100, oracle/build_ref_gpu.py)"""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(name):
    import torch  # noqa: F401  libtorch first
    full = f"refgpu_{name}"
    if full in sys.modules:
        return sys.modules[full]
    path = os.path.join(ROOT, "oracle", "_ref", full + ".so")
    if not os.path.exists(path):
        raise ImportError(f"{path} missing: build it with `python oracle/build_ref_gpu.py` in the authoring container")
    spec = importlib.util.spec_from_file_location(full, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules[full] = mod
    return mod
