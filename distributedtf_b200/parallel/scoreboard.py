"""Device-side eval-score board.

Reference behaviour: workers pickle ``[[id, accuracy, hparams], ...]`` and the master receives
them one rank after another (``pbt_cluster.py:114-128``).  Here the evaluation kernels count
hits on the device; ``publish`` turns the count into an accuracy and stores it, followed by a
release-store of the round epoch, into the member's slot of a board that lives in memory the
master can see (a peer pointer when master and member sit on different GPUs).  The master
polls the board until every live member shows the round's epoch, then sorts at most a few
dozen floats.  The control-plane GET remains the carrier of the hyper-parameter dicts.
"""
from __future__ import annotations

import ctypes as C
import time
from typing import Dict, Iterable, Optional

import torch

from ..ops import native, native_ext
from . import symm

vp, ull, i32 = C.c_void_p, C.c_ulonglong, C.c_int
native_ext.register("dtf_score_publish", [vp, vp, i32, ull, i32, vp, vp])
SLOT_BYTES = 16


class ScoreBoard:
    def __init__(self, max_members: int, device):
        self.max_members = max_members
        self.buf = symm.DeviceBuffer(max_members * SLOT_BYTES, device)
        self.device = self.buf.device
        self._host = torch.zeros(max_members * SLOT_BYTES, dtype=torch.uint8,
                                 pin_memory=torch.cuda.is_available())
        self._stream = torch.cuda.Stream(device=self.device)

    @property
    def ptr(self) -> int:
        return self.buf.ptr

    def handle(self) -> bytes:
        return self.buf.ipc_handle()

    def read(self, epoch: int, members: Optional[Iterable[int]] = None, timeout: float = 60.0) -> Dict[int, float]:
        """Block until every listed member published ``epoch`` (>=); returns {member: score}."""
        want = list(members) if members is not None else None
        deadline = time.time() + timeout
        while True:
            with torch.cuda.stream(self._stream):
                self._host.copy_(self.buf.tensor[: self._host.numel()], non_blocking=True)
            self._stream.synchronize()
            scores = self._host.view(torch.float32).view(-1, 4)[:, 0]
            epochs = self._host.view(torch.int64).view(-1, 2)[:, 1]
            ids = want if want is not None else [i for i in range(self.max_members) if int(epochs[i]) > 0]
            if all(int(epochs[i]) >= epoch for i in ids):
                return {i: float(scores[i]) for i in ids}
            if time.time() > deadline:
                raise TimeoutError(f"score board: members {[i for i in ids if int(epochs[i]) < epoch]} "
                                   f"did not publish epoch {epoch}")
            time.sleep(0.0002)


def publish(board_ptr: int, member_id: int, correct_cell: torch.Tensor, total: int, epoch: int,
            reset: bool = True) -> None:
    """Launch the publish kernel on the current stream; ``board_ptr`` may be a peer pointer."""
    slot = board_ptr + member_id * SLOT_BYTES
    native.check(native.lib().dtf_score_publish(vp(slot), vp(correct_cell.data_ptr()), int(total), int(epoch),
                                                int(reset), vp(correct_cell.data_ptr()), native.stream_ptr()),
                 "dtf_score_publish")
