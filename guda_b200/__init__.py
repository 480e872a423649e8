"""guda_b200 -- B200-native (sm_100a) replacement for the BLAS hot path of LynnColeArt/guda.

Layout:
  csrc/     hand-written CUDA kernels + the C ABI (include/guda_b200.h) -> lib/libguda_b200.so
  capi.py   ctypes binding of that C ABI (what the Go cgo shim in go/guda binds)
  api.py    host-side mirror of GUDA's root-package API (Malloc/Memcpy/AXPY/DOT/GEMM/Fused*...)
There is no CPU fallback: importing works anywhere, computing needs a B200.
"""
from .api import *          # noqa: F401,F403
from .api import DevicePtr, Stream, Event, ErrNotSupported   # noqa: F401
from . import capi  # noqa: F401
from .capi import BlasPanic, GUDAError, LIB_PATH, load       # noqa: F401

__version__ = "0.1.0"
