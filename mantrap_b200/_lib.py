"""Loader of the CUDA C-ABI library. There is NO fallback: a missing library or a missing GPU is an error."""
import ctypes
import os

from . import _cabi

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libmantrap_b200.so")
_LIB = None


def load() -> ctypes.CDLL:
    """Load (once) and bind `libmantrap_b200.so`; raises if it has not been built (`python -m mantrap_b200.build`)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m mantrap_b200.build` "
                               "(mantrap_b200 has no CPU fallback)")
        lib = _cabi.bind(ctypes.CDLL(LIB_PATH))
        if lib.mtb_abi_version() != _cabi.ABI_VERSION:
            raise RuntimeError("mantrap_b200: ABI version mismatch between the Python package and the library")
        _LIB = lib
    return _LIB
