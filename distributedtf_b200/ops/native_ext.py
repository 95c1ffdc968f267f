"""Late-bound declarations for entry points that live in later source files (exploit transfer,
fused all-reduce + update, score board, symmetric memory, host IO).  A module that wraps such
an entry point calls ``register`` at import time; the signature is applied as soon as the
library is (or already has been) loaded."""
from __future__ import annotations

import ctypes as C
from typing import Optional

_SIGS = {}
_LIB: Optional[C.CDLL] = None


def _apply(l: C.CDLL, name: str) -> None:
    argtypes, restype = _SIGS[name]
    fn = getattr(l, name, None)
    if fn is None:
        raise AttributeError(f"libdtf_b200.so does not export {name}: rebuild with build_native.py")
    fn.argtypes = argtypes
    fn.restype = restype


def register(name: str, argtypes, restype=C.c_int) -> None:
    _SIGS[name] = (argtypes, restype)
    if _LIB is not None:
        _apply(_LIB, name)


def declare(l: C.CDLL) -> None:
    global _LIB
    _LIB = l
    for name in _SIGS:
        _apply(l, name)
