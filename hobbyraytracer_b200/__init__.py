"""hobbyraytracer_b200 — B200-native path-tracing core behind the host surface of Todegal/HobbyRaytracer.

The product is the CUDA shared library `csrc/libhrt_b200.so` (C ABI: include/hrt.h).  This package is the thin
Python host side used by tests and benchmarks: a ctypes binding (`abi`), procedural stand-ins for the benchmark
scenes (`scenes`) and the multi-GPU driver (`multigpu`).  There is no CPU fallback: `load()` raises if the
CUDA library has not been built.
"""
import os

from . import abi

PRODUCT_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libhrt_b200.so")

_lib = None


def load() -> "abi.Lib":
    """Bind csrc/libhrt_b200.so.  Raises HrtError when the library is missing (build with __graft_entry__.build())."""
    global _lib
    if _lib is None:
        so = os.environ.get("HRT_PRODUCT_SO", PRODUCT_SO)      # A/B builds of the same CUDA library (tuning experiments)
        if so != PRODUCT_SO:
            _lib = abi.Lib(so, "hrt_")
            return _lib
        if not os.path.exists(PRODUCT_SO):
            raise abi.HrtError(f"{PRODUCT_SO} is not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        _lib = abi.Lib(PRODUCT_SO, "hrt_")
    return _lib
