"""paddle-lite_b200 -- B200 (sm_100a) int8 conv/GEMM hot path behind Paddle-Lite's
kCUDA/kInt8 kernel plugin surface.

The product is the C-ABI shared library `libpl_int8_b200.so` (include/pl_int8.h)
plus the C++ mirror of the reference's kernel-plugin interface under `lite/`.
This Python package is only the loader used by tests/, bench.py and
__graft_entry__.py (the directory name has a hyphen, so it is imported through
importlib -- see `load()` in tests/conftest.py / __graft_entry__.py).
There is no CPU fallback: importing `capi` without the built library raises.
"""
from . import capi  # noqa: F401
from . import lite_harness  # noqa: F401
from . import nsplit  # noqa: F401
