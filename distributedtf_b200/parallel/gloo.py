"""Control plane over ``torch.distributed`` (gloo): one endpoint per process.

Used when the job is launched with one *extra* process for the master
(reference topology: rank 0 never trains, ``main_manager.py:46-73``).  Objects
are pickled into uint8 tensors; a length tensor precedes the payload.  Only CPU
tensors travel here — the tensor path never uses this transport.
"""
from __future__ import annotations

import time
from typing import Any, Optional

import torch
import torch.distributed as dist

from .comm_base import Comm, CommTimeout, Request, dumps, loads


class GlooComm(Comm):
    def __init__(self, group: Optional["dist.ProcessGroup"] = None):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised (init_process_group('gloo'))")
        self._group = group
        self._rank = dist.get_rank(group)
        self._size = dist.get_world_size(group)

    def Get_rank(self) -> int:
        return self._rank

    def Get_size(self) -> int:
        return self._size

    def isend(self, obj: Any, dest: int) -> Request:
        payload = torch.frombuffer(bytearray(dumps(obj)), dtype=torch.uint8)
        header = torch.tensor([payload.numel()], dtype=torch.int64)
        works = [dist.isend(header, dest, group=self._group),
                 dist.isend(payload, dest, group=self._group)]
        keep = (header, payload)

        def waiter(works=works, keep=keep):
            for w in works:
                w.wait()

        return Request(waiter)

    def _recv_tensor(self, t: torch.Tensor, source: int, deadline: Optional[float]) -> None:
        if deadline is None:
            dist.recv(t, source, group=self._group)
            return
        # a dead worker must not hang the master: poll the non-blocking receive
        work = dist.irecv(t, source, group=self._group)
        while not work.is_completed():
            if time.monotonic() > deadline:
                raise CommTimeout(f"gloo control plane: no message from endpoint {source} within the timeout")
            time.sleep(0.002)
        work.wait()

    def recv(self, source: int, timeout: Optional[float] = None) -> Any:
        deadline = None if timeout is None else time.monotonic() + float(timeout)
        header = torch.zeros(1, dtype=torch.int64)
        self._recv_tensor(header, source, deadline)
        payload = torch.empty(int(header.item()), dtype=torch.uint8)
        self._recv_tensor(payload, source, deadline)
        return loads(payload.numpy().tobytes())
