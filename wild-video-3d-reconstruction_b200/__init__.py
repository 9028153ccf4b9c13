"""B200-native (sm_100a) altcorr + fastba for the DPVO tracker of Wild-video-3d-reconstruction.

Importing this package puts its directory on sys.path so that the reference's unchanged host code finds the
two extension modules under the names it imports (`import cuda_corr`, dpvo/altcorr/correlation.py:2;
`import cuda_ba`, dpvo/fastba/ba.py:2).  There is no CPU or PyTorch fallback: if the native modules are
missing the import fails, and on a machine with a GPU that is not sm_100 the first call fails.

    pkg = importlib.import_module("wild-video-3d-reconstruction_b200")
    pkg.altcorr.corr(...), pkg.fastba.BA(...)          # mirrors of dpvo.altcorr / dpvo.fastba
    pkg.install_as_dpvo_backends()                      # make `import cuda_corr, cuda_ba` resolve to this build
"""
import ctypes
import importlib
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libwvb200.so")


def install_as_dpvo_backends():
    """make `import cuda_corr` / `import cuda_ba` resolve to this build (idempotent)"""
    if HERE not in sys.path:
        sys.path.insert(0, HERE)


def native_modules():
    """(cuda_corr, cuda_ba) -- raises ImportError with the build hint when they were never built"""
    install_as_dpvo_backends()
    import torch  # noqa: F401  libtorch must be loaded before the extensions
    try:
        return importlib.import_module("cuda_corr"), importlib.import_module("cuda_ba")
    except ImportError as e:  # fail loudly: no fallback path exists
        raise ImportError(f"native modules missing ({e}); build them with `python {HERE}/build.py`") from e


def native_agg():
    """cuda_agg: group plan + fused scatter-softmax / scatter-sum (the reference calls torch_scatter here, dpvo/blocks.py:41-43)"""
    install_as_dpvo_backends()
    import torch  # noqa: F401
    try:
        return importlib.import_module("cuda_agg")
    except ImportError as e:
        raise ImportError(f"native module cuda_agg missing ({e}); build it with `python {HERE}/build.py`") from e


def capi():
    """ctypes handle on the C ABI (include/wvb200.h)"""
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} missing; build it with `python {HERE}/build.py`")
    return ctypes.CDLL(LIB_PATH)


def __getattr__(name):
    if name in ("altcorr", "fastba", "synthetic", "hostapi", "dist", "projective_ops", "blocks", "colmap_export"):
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
