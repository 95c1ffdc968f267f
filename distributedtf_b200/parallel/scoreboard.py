"""Device-side eval-score board.

Reference behaviour: workers pickle ``[[id, accuracy, hparams], ...]`` and the master receives
them one rank after another (``pbt_cluster.py:114-128``); that gather is also the round barrier
(``:191-193``).  Here every member owns a 16-byte slot ``{hits u32, seen u32, epoch u64}`` of a board
that lives in memory the master can see (a peer pointer when master and member sit on different
GPUs).  A member's evaluation kernels add their hits and batch sizes straight into the slot with
system-scope reductions (``DtfHead.board_hits``); after the last batch one tiny kernel release-stores
the round's epoch.  No host is involved on the member's side, so the score is on the board as soon
as the evaluation kernels have run — before the member reads its own accuracy back, writes its CSV
row or hands its checkpoint to the writer.  The master polls the board until every live member shows
the round's epoch, then sorts at most a few dozen numbers; with the board live it no longer waits for
GET replies to rank (``master.PBTCluster.exploit``).  A member whose loss went non-finite, or whose
training raised, shows up as a *dropped* marker, so the master never waits for a member its worker
has removed.
"""
from __future__ import annotations

import ctypes as C
import math
import time
from typing import Dict, Iterable, Optional

import torch

from ..ops import native, native_ext
from . import symm
from ..utils.streams import new_stream

vp, ull, i32 = C.c_void_p, C.c_ulonglong, C.c_int
native_ext.register("dtf_score_publish", [vp, vp, i32, ull, i32, vp, vp])
native_ext.register("dtf_score_reset", [vp, vp])
native_ext.register("dtf_score_release", [vp, ull, vp, i32, vp])
SLOT_BYTES = 16
DROPPED = 0xFFFFFFFF


class ScoreBoard:
    """The board, or a view of it.  ``ScoreBoard(n, device)`` allocates (the master's side);
    ``ScoreBoard.view(ptr, n, device)`` wraps an already mapped (peer) pointer for reading."""

    def __init__(self, max_members: int, device, _ptr: Optional[int] = None):
        self.max_members = max_members
        self.device = torch.device(device)
        if _ptr is None:
            self.buf = symm.DeviceBuffer(max_members * SLOT_BYTES, self.device)
            self._ptr = self.buf.ptr
            self._tensor = self.buf.tensor
        else:
            self.buf = None
            self._ptr = int(_ptr)
            self._owner = symm._RawCuda(self._ptr, max_members * SLOT_BYTES)
            with torch.cuda.device(self.device):
                self._tensor = torch.as_tensor(self._owner, device=self.device)
        self._host = torch.zeros(max_members * SLOT_BYTES, dtype=torch.uint8,
                                 pin_memory=torch.cuda.is_available())
        self._stream = new_stream(self.device)

    @classmethod
    def view(cls, ptr: int, max_members: int, device) -> "ScoreBoard":
        return cls(max_members, device, _ptr=ptr)

    @property
    def ptr(self) -> int:
        return self._ptr

    def handle(self) -> bytes:
        return self.buf.ipc_handle()

    def snapshot(self):
        """One copy of the board to the host: (hits[int], seen[int], epochs[int]) lists."""
        from ..utils.capture_gate import GATE
        GATE.enter_step()                 # never while a thread of this process captures a CUDA graph
        try:
            with torch.cuda.stream(self._stream):
                self._host.copy_(self._tensor[: self._host.numel()], non_blocking=True)
            self._stream.synchronize()
        finally:
            GATE.exit_step()
        words = self._host.view(torch.int32).view(-1, 4)
        epochs = self._host.view(torch.int64).view(-1, 2)[:, 1]
        return ([int(w) & 0xFFFFFFFF for w in words[:, 0]], [int(w) & 0xFFFFFFFF for w in words[:, 1]],
                [int(e) for e in epochs])

    def read(self, epoch: int, members: Optional[Iterable[int]] = None, timeout: float = 60.0,
             poll: float = 0.0001) -> Dict[int, float]:
        """Block until every listed member published ``epoch`` (>=); returns {member: score}, NaN for a member
        that was dropped."""
        want = list(members) if members is not None else None
        deadline = time.time() + timeout
        while True:
            hits, seen, epochs = self.snapshot()
            ids = want if want is not None else [i for i in range(self.max_members) if epochs[i] > 0]
            if all(epochs[i] >= epoch for i in ids):
                return {i: (float("nan") if hits[i] == DROPPED else hits[i] / max(seen[i], 1)) for i in ids}
            if time.time() > deadline:
                raise TimeoutError(f"score board: members {[i for i in ids if epochs[i] < epoch]} "
                                   f"did not publish epoch {epoch}")
            if poll > 0:
                time.sleep(poll)


def slot_ptr(board_ptr: int, member_id: int) -> int:
    return int(board_ptr) + int(member_id) * SLOT_BYTES


def publish(board_ptr: int, member_id: int, correct_cell: torch.Tensor, total: int, epoch: int,
            reset: bool = True) -> None:
    """Store (hits, seen) and release ``epoch`` in one launch on the current stream (data-parallel members:
    the replicas' counts were summed first); ``board_ptr`` may be a peer pointer."""
    native.check(native.lib().dtf_score_publish(vp(slot_ptr(board_ptr, member_id)), vp(correct_cell.data_ptr()),
                                                int(total), int(epoch), int(reset), vp(correct_cell.data_ptr()),
                                                native.stream_ptr()), "dtf_score_publish")


def reset(board_ptr: int, member_id: int) -> None:
    """Start of an evaluation: zero the slot's counters (current stream)."""
    native.check(native.lib().dtf_score_reset(vp(slot_ptr(board_ptr, member_id)), native.stream_ptr()),
                 "dtf_score_reset")


def release(board_ptr: int, member_id: int, epoch: int, loss: Optional[torch.Tensor] = None,
            dropped: bool = False) -> None:
    """End of an evaluation: release-store ``epoch`` behind the evaluation kernels of the current stream; a
    non-finite ``loss`` (device scalar) or ``dropped`` turns the entry into the dropped marker."""
    native.check(native.lib().dtf_score_release(vp(slot_ptr(board_ptr, member_id)), int(epoch),
                                                vp(loss.data_ptr() if loss is not None else 0), int(dropped),
                                                native.stream_ptr()), "dtf_score_release")


def is_nan(x: float) -> bool:
    return isinstance(x, float) and math.isnan(x)
