"""Control plane over ``torch.distributed`` (gloo): one endpoint per process.

Used when the job is launched with one *extra* process for the master
(reference topology: rank 0 never trains, ``main_manager.py:46-73``).  Objects
are pickled into uint8 tensors; a length tensor precedes the payload.  Only CPU
tensors travel here — the tensor path never uses this transport.
"""
from __future__ import annotations

from typing import Any, Optional

import torch
import torch.distributed as dist

from .comm_base import Comm, Request, dumps, loads


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

    def recv(self, source: int, timeout: Optional[float] = None) -> Any:
        header = torch.zeros(1, dtype=torch.int64)
        dist.recv(header, source, group=self._group)
        payload = torch.empty(int(header.item()), dtype=torch.uint8)
        dist.recv(payload, source, group=self._group)
        return loads(payload.numpy().tobytes())
