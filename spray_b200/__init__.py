"""spray_b200 — B200-native (sm_100a CUDA) implementation of SpRay's per-domain ray-tracing hot path.

The product is libspraygpu.so (spray_b200/csrc, C ABI in include/spray_gpu.h) plus the C++ host mirror
of the reference's plugin seams (host/).  This Python package is the thin harness used by tests and
bench.py: a ctypes binding of the C ABI (capi), scene/camera helpers (scenes) and the multi-GPU
domain-sharded frame loop over torch.distributed (multigpu).
"""
from . import scenes  # noqa: F401

__all__ = ["scenes", "capi"]
