"""guda_b200 -- B200-native (sm_100a) replacement for the BLAS hot path of LynnColeArt/guda.

Layout:
  csrc/     hand-written CUDA kernels + the C ABI (include/guda_b200.h) -> lib/libguda_b200.so
  capi.py   ctypes binding of that C ABI (what the Go cgo shim in go/guda binds)
  api.py    host-side mirror of GUDA's root-package API (Malloc/Memcpy/AXPY/DOT/GEMM/Fused*...)
  tolerance.py  the architecture-aware tolerance framework (tolerance.go, tolerance_arch.go) with the sm_100a entry
There is no CPU fallback: importing works anywhere, computing needs a B200.
"""
from .api import *          # noqa: F401,F403
from .api import DevicePtr, Stream, Event, ErrNotSupported   # noqa: F401
from . import capi  # noqa: F401
from . import tolerance  # noqa: F401
from .tolerance import (ArchToleranceConfig, DefaultTolerance, Float32NearEqual, Float32ULPDiff, GetArchTolerance,      # noqa: F401
                        GetOperationTolerance, IsARM64, OperationTolerance, RelaxedTolerance, StrictTolerance, ToleranceConfig,
                        VerificationResult, VerifyFloat32Array, mergeTolerances)
from .capi import BlasPanic, GUDAError, LIB_PATH, load       # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):      # GEMMArchTolerance, ConvArchTolerance, ReduceArchTolerance, StandardTolerances (read from the C library on use)
    if name in ("GEMMArchTolerance", "ConvArchTolerance", "ReduceArchTolerance", "StandardTolerances"):
        return getattr(tolerance, name)
    raise AttributeError(name)
