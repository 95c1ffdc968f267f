"""Declarations for entry points that live in later source files (exploit transfer,
fused all-reduce + update, score board, symmetric memory, host IO).  Kept separate so the
core bindings stay readable; every symbol is optional at declaration time and checked at
call time."""
from __future__ import annotations

import ctypes as C

_SIGS = {}


def register(name: str, argtypes, restype=C.c_int) -> None:
    _SIGS[name] = (argtypes, restype)


def declare(l: C.CDLL) -> None:
    for name, (argtypes, restype) in _SIGS.items():
        fn = getattr(l, name, None)
        if fn is not None:
            fn.argtypes = argtypes
            fn.restype = restype
