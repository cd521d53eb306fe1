"""Loader of libstanmath_cuda.so (the CUDA kernels + C ABI).  No fallback: if the library is
missing it is built with nvcc; if that fails, or no B200 is present at call time, calls raise."""
import ctypes
import os

from . import _abi
from .build import LIB, build

_lib = None


class StanCudaError(RuntimeError):
    """CUDA / NCCL / argument failure reported by the C ABI (negative return code)."""

    def __init__(self, rc, msg):
        super().__init__("stan_cuda error %d: %s" % (rc, msg))
        self.rc = rc


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = _abi.bind(ctypes.CDLL(LIB))
    return _lib


def check(rc):
    """Raise on failure codes; pass OK / NONFINITE through."""
    if rc < 0:
        raise StanCudaError(rc, lib().stan_cuda_last_error().decode())
    return rc
