"""In-process control-plane fabric: one queue per (source, dest) pair.

Fills the reference's testing gap (no fake communicator
exists there): the master and any number of workers run as threads of one
process and talk through ``LocalFabric.endpoint(rank)`` objects that satisfy the
``Comm`` contract, including pickling in transit so no object is shared.
"""
from __future__ import annotations

import queue
import threading
import time
from typing import Any, Dict, Optional, Tuple

from .comm_base import Comm, CommTimeout, Request, dumps, loads


class LocalFabric:
    def __init__(self, size: int, delay_fn=None):
        if size < 2:
            raise ValueError("a PBT fabric needs a master and at least one worker (size >= 2)")
        self.size = size
        self._queues: Dict[Tuple[int, int], "queue.Queue[bytes]"] = {
            (s, d): queue.Queue() for s in range(size) for d in range(size)
        }
        self._delay_fn = delay_fn  # optional fault/latency injection: f(src, dst) -> seconds
        self.sent = 0
        self._lock = threading.Lock()
        self._aborted: Optional[BaseException] = None

    def abort(self, exc: BaseException) -> None:
        """An endpoint's thread died: every blocked or future ``recv`` raises instead of waiting forever."""
        self._aborted = exc

    def endpoint(self, rank: int) -> "LocalComm":
        if not 0 <= rank < self.size:
            raise ValueError(f"rank {rank} outside fabric of size {self.size}")
        return LocalComm(self, rank)

    def _put(self, src: int, dst: int, payload: bytes) -> None:
        if self._delay_fn is not None:
            d = self._delay_fn(src, dst)
            if d:
                time.sleep(d)
        with self._lock:
            self.sent += 1
        self._queues[(src, dst)].put(payload)

    def _get(self, src: int, dst: int, timeout: Optional[float]) -> bytes:
        deadline = None if timeout is None else time.monotonic() + timeout
        q = self._queues[(src, dst)]
        while True:                                # short polls so that abort() is noticed
            if self._aborted is not None:
                raise RuntimeError(f"fabric aborted: {self._aborted!r}") from self._aborted
            left = 0.25 if deadline is None else min(0.25, deadline - time.monotonic())
            if left <= 0:
                raise CommTimeout(f"endpoint {dst}: no message from {src} within {timeout}s")
            try:
                return q.get(timeout=left)
            except queue.Empty:
                continue


class LocalComm(Comm):
    def __init__(self, fabric: LocalFabric, rank: int):
        self._fabric = fabric
        self._rank = rank

    def Get_rank(self) -> int:
        return self._rank

    def Get_size(self) -> int:
        return self._fabric.size

    def isend(self, obj: Any, dest: int) -> Request:
        self._fabric._put(self._rank, dest, dumps(obj))
        return Request()

    def recv(self, source: int, timeout: Optional[float] = None) -> Any:
        return loads(self._fabric._get(source, self._rank, timeout))
