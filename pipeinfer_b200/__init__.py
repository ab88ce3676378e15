"""pipeinfer_b200 -- B200-native implementation of PipeInfer's asynchronous pipelined speculation hot path.

The product is `libpipeinfer_b200.so` (hand-written sm_100a CUDA + C++ host runtime behind a C ABI, see include/*.h).
This Python package is only the thin ctypes mirror used by tests and bench.py; there is no CPU fallback: importing
`pipeinfer_b200.lib` raises if the shared library has not been built (python __graft_entry__.py build).
"""
from .lib import Lib, load, LibraryMissing  # noqa: F401

__all__ = ["Lib", "load", "LibraryMissing"]
