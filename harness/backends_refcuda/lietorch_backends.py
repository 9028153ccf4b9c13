"""`import lietorch_backends` for the REF-CUDA arm of the harness: re-exports the reference's own lietorch_backends built for sm_100 (oracle/_ref/refgpu_lietorch_backends.so)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _refgpu  # noqa: E402

_m = _refgpu.load("lietorch_backends")
globals().update({k: getattr(_m, k) for k in dir(_m) if not k.startswith("__")})
