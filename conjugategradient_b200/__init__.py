"""B200-native Conjugate Gradient hot path (CUDA, sm_100a) behind the reference's API.

The product is the C-ABI library ``libcgb200.so`` (include/cgb200.h) and the drop-in C++ headers
in ``include/``.  This package holds the CUDA sources (``csrc/``), the ctypes binding and a thin
Python harness that mirrors the reference's class names for tests and the benchmark.
"""
from . import generators  # noqa: F401  (NumPy only; no GPU needed)

__all__ = ["generators", "solver", "_lib"]
