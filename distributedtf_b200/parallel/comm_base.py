"""Endpoint API of the control plane.

Parity: the subset of ``mpi4py.MPI.Comm`` the reference uses directly
(``pbt_cluster.py:64-77,93-98,114-128``; ``training_worker.py:26,38,45``):
``Get_rank, Get_size, isend(obj, dest) -> req; req.wait(), send(obj, dest),
recv(source)``.  Guarantees every transport must give:

* per (source, dest) pair messages are delivered in send order (MPI
  non-overtaking), so each worker sees a FIFO instruction queue;
* ``recv(source=i)`` only returns messages from endpoint ``i``;
* objects are deep-copied in transit (pickle), never shared.

Extension over MPI: ``recv(source, timeout=seconds)`` raises ``CommTimeout`` so a
dead worker cannot block the master forever.
"""
from __future__ import annotations

import pickle
from typing import Any, Optional


class CommTimeout(TimeoutError):
    pass


class Request:
    """Completed-or-pending send handle (``mpi4py`` ``Request`` look-alike)."""

    def __init__(self, waiter=None):
        self._waiter = waiter

    def wait(self) -> None:
        if self._waiter is not None:
            self._waiter()
            self._waiter = None

    Wait = wait


class Comm:
    def Get_rank(self) -> int:  # noqa: N802 (mpi4py naming)
        raise NotImplementedError

    def Get_size(self) -> int:  # noqa: N802
        raise NotImplementedError

    def isend(self, obj: Any, dest: int) -> Request:
        raise NotImplementedError

    def send(self, obj: Any, dest: int) -> None:
        self.isend(obj, dest).wait()

    def recv(self, source: int, timeout: Optional[float] = None) -> Any:
        raise NotImplementedError

    def close(self) -> None:
        pass

    # lower-case aliases
    @property
    def rank(self) -> int:
        return self.Get_rank()

    @property
    def size(self) -> int:
        return self.Get_size()


def dumps(obj: Any) -> bytes:
    return pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)


def loads(buf: bytes) -> Any:
    return pickle.loads(buf)
