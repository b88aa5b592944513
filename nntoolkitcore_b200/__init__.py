"""nntoolkitcore_b200 -- B200-native (sm_100a) implementation of NNToolkitCore's batched inference forward path.

The product is the C library `lib/libnntoolkitcore_b200.so` (C-ABI declared in include/nntoolkitcore/*.h and
include/nntk_b200_shim.h).  This package only builds it (`build()`), dlopens it (`lib()`) and offers a thin ctypes
mirror of the reference's create/apply API (`layers`).  There is no CPU fallback: without the built library `lib()`
raises, and without a CUDA device every Apply fails with NNTK_ERR_CUDA.
"""
from __future__ import annotations

from pathlib import Path

from . import capi

LIB_PATH = Path(__file__).resolve().parent / "lib" / "libnntoolkitcore_b200.so"
_lib = None


def build(force: bool = False):
    from .build import build as _build
    return _build(force=force)


def lib():
    """The bound product library.  Raises if it has not been built -- never substitutes anything else."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} is missing: run `python -m nntoolkitcore_b200.build` (needs nvcc). "
                              "There is no CPU fallback.")
        _lib = capi.bind(str(LIB_PATH), ext=True)
    return _lib
