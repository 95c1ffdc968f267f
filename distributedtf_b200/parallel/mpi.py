"""Optional mpi4py adapter.

When the job runs under ``mpirun`` and mpi4py imports, ``MPI.COMM_WORLD`` already
has the ``Comm`` API (that is where the API comes from: reference
``main_manager.py:46``).  Neither is installed on the target image, so this is a
thin, gated wrapper that only adds the ``timeout`` keyword.
"""
from __future__ import annotations

import time
from typing import Any, Optional

from .comm_base import Comm, CommTimeout, Request


def available() -> bool:
    try:
        import mpi4py  # noqa: F401
        return True
    except Exception:
        return False


class MpiComm(Comm):
    def __init__(self):
        from mpi4py import MPI  # gated import
        self._MPI = MPI
        self._comm = MPI.COMM_WORLD

    def Get_rank(self) -> int:
        return self._comm.Get_rank()

    def Get_size(self) -> int:
        return self._comm.Get_size()

    def isend(self, obj: Any, dest: int) -> Request:
        req = self._comm.isend(obj, dest=dest)
        return Request(req.wait)

    def recv(self, source: int, timeout: Optional[float] = None) -> Any:
        if timeout is None:
            return self._comm.recv(source=source)
        deadline = time.time() + timeout
        while not self._comm.iprobe(source=source):
            if time.time() > deadline:
                raise CommTimeout(f"no MPI message from {source} within {timeout}s")
            time.sleep(0.001)
        return self._comm.recv(source=source)
