"""quest-b200: a B200-native (sm_100a) gate-application engine behind the QuEST C API.

The product is the shared library ``quest_b200/lib/libQuEST.so`` (hand-written CUDA kernels behind the
unchanged ``QuEST.h`` entry points).  This package holds its sources (``csrc/``), the build recipe
(``build.py``) and a ctypes mirror of the C API (``capi``) used by the tests and ``bench.py``.
"""
from . import capi  # noqa: F401

__all__ = ["capi"]
