import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

PKG = "wild-video-3d-reconstruction_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def syn():
    return importlib.import_module(PKG + ".synthetic")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    d = os.path.join(ROOT, "tests", "golden")
    return {k: np.load(os.path.join(d, k + ".npz")) for k in ("corr_ref", "ba_ref", "pyba_ref", "pops_ref", "ba_loop_ref")}


@pytest.fixture(scope="session")
def ref_modules():
    """the reference's own sources built as host code (oracle/_ref); None when not built"""
    import build_ref
    return {n: build_ref.load(n) for n in ("ref_cuda_corr", "ref_cuda_ba", "ref_lietorch")}
