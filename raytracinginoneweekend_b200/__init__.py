"""B200-native path-tracing backend for the Ray Tracing in One Weekend renderer.

The product is ``librtiow_b200.so`` (hand-written sm_100a CUDA behind the C ABI of include/rt_api.h);
this package is its Python face: ``backend`` (ctypes binding), ``scenes`` (scene definitions) and
``multigpu`` (one process per GPU, tile stealing, NCCL gather).
"""
from . import scenes  # noqa: F401

__all__ = ["scenes"]
