"""pint_b200 -- B200-native batch engine for pint packed integers.

The product is libpint_cuda.so (hand-written sm_100a kernels behind the C ABI of
include/pint_cuda.h); `include/pint/cuda.hpp` is its C++ template layer.  This
package is the Python mirror of that layer, used by the tests and bench.py:
`Layout` == packed_int<Integer, Bits...>, `ops` == the batched operations.
"""
from . import _lib  # noqa: F401
from .layout import Layout, storage_bytes  # noqa: F401

__all__ = ["Layout", "storage_bytes", "ops", "sharded", "dense"]


def __getattr__(name):
    if name in ("ops", "sharded", "dense"):
        import importlib
        return importlib.import_module(f".{name}", __name__)
    raise AttributeError(name)
