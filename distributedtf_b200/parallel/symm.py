"""Symmetric device memory: buffers that every rank of the node can address.

SURVEY §5.8 data plane.  Each population member owns one device buffer
``[ state | snapshot 0 | snapshot 1 | control ]`` allocated with ``cudaMalloc`` through the
native library (not the torch caching allocator, so the allocation can be exported).  Its
CUDA IPC handle is published to the other worker processes once, after ``ADD_GRAPHS``; a
peer maps it lazily on first use and from then on reads / writes it directly from inside
kernels over NVLink (``csrc/comm_kernels.cu``).  Members living in the same process skip
IPC and use the raw pointer.

The handle exchange itself is control-plane traffic (64-byte opaque handles) and goes over
``torch.distributed`` object collectives — never tensors.
"""
from __future__ import annotations

import ctypes as C
import threading
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

from ..ops import native, native_ext

vp, ll, i32 = C.c_void_p, C.c_longlong, C.c_int
native_ext.register("dtf_symm_alloc", [ll, C.POINTER(vp)])
native_ext.register("dtf_symm_free", [vp])
native_ext.register("dtf_ipc_get_handle", [vp, C.c_char_p])
native_ext.register("dtf_ipc_open_handle", [C.c_char_p, C.POINTER(vp)])
native_ext.register("dtf_ipc_close", [vp])
native_ext.register("dtf_enable_peer", [i32])
native_ext.register("dtf_last_cuda_error", [C.c_char_p, i32])


class _RawCuda:
    """Minimal ``__cuda_array_interface__`` carrier so torch can view a raw allocation."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3, "strides": None}


class DeviceBuffer:
    """An exportable device allocation viewed as a torch uint8 tensor."""

    def __init__(self, nbytes: int, device: torch.device):
        self.nbytes = (nbytes + 255) // 256 * 256
        self.device = torch.device(device)
        with torch.cuda.device(self.device):
            out = vp()
            native.check(native.lib().dtf_symm_alloc(self.nbytes, C.byref(out)), "dtf_symm_alloc")
            self.ptr = int(out.value)
            self._owner = _RawCuda(self.ptr, self.nbytes)
            self.tensor = torch.as_tensor(self._owner, device=self.device)
        self._freed = False

    def ipc_handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        with torch.cuda.device(self.device):
            native.check(native.lib().dtf_ipc_get_handle(vp(self.ptr), buf), "dtf_ipc_get_handle")
        return buf.raw

    def free(self) -> None:
        if not self._freed:
            self._freed = True
            self.tensor = None
            with torch.cuda.device(self.device):
                native.lib().dtf_symm_free(vp(self.ptr))

    def __call__(self, nbytes: int) -> torch.Tensor:
        """``arena_alloc`` callback shape: returns the first ``nbytes`` of the buffer."""
        if nbytes > self.nbytes:
            raise ValueError("symmetric buffer too small")
        return self.tensor[:nbytes]


@dataclass
class Entry:
    rank: int            # worker process index that owns the member
    device: int
    nbytes: int
    handle: Optional[bytes]
    local_ptr: Optional[int] = None


class Directory:
    """cluster_id -> how to reach that member's buffer from this process."""

    def __init__(self):
        self.entries: Dict[int, Entry] = {}
        self._opened: Dict[int, int] = {}
        self._lock = threading.Lock()
        self.my_rank = 0

    def register_local(self, cluster_id: int, buf: DeviceBuffer, rank: int) -> None:
        with self._lock:
            self.entries[cluster_id] = Entry(rank, buf.device.index or 0, buf.nbytes, None, buf.ptr)

    def export_local(self, buffers: Dict[int, DeviceBuffer], rank: int) -> List[Tuple[int, int, int, int, bytes]]:
        return [(cid, rank, b.device.index or 0, b.nbytes, b.ipc_handle()) for cid, b in buffers.items()]

    def import_remote(self, records) -> None:
        with self._lock:
            for cid, rank, dev, nbytes, handle in records:
                if cid in self.entries and self.entries[cid].local_ptr is not None:
                    continue
                self.entries[cid] = Entry(rank, dev, nbytes, handle)

    def pointer(self, cluster_id: int) -> int:
        """Device address of member ``cluster_id``'s buffer, valid in the current process."""
        with self._lock:
            e = self.entries.get(cluster_id)
            if e is None:
                raise KeyError(f"member {cluster_id} is not in the symmetric directory")
            if e.local_ptr is not None:
                return e.local_ptr
            if cluster_id not in self._opened:
                out = vp()
                rc = native.lib().dtf_ipc_open_handle(e.handle, C.byref(out))
                if rc != 0:
                    msg = C.create_string_buffer(256)
                    native.lib().dtf_last_cuda_error(msg, 256)
                    raise native.NativeError(f"cudaIpcOpenMemHandle failed for member {cluster_id}: {msg.value.decode()}")
                self._opened[cluster_id] = int(out.value)
            return self._opened[cluster_id]

    def close(self) -> None:
        with self._lock:
            for p in self._opened.values():
                native.lib().dtf_ipc_close(vp(p))
            self._opened.clear()


DIRECTORY = Directory()


def exchange_directory(local_buffers: Dict[int, DeviceBuffer], group=None) -> None:
    """Collective over the worker group: publish my members' handles, learn everyone else's."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    rank = dist.get_rank(group)
    DIRECTORY.my_rank = rank
    mine = DIRECTORY.export_local(local_buffers, rank)
    gathered: List[Optional[list]] = [None] * dist.get_world_size(group)
    dist.all_gather_object(gathered, mine, group=group)
    for r, recs in enumerate(gathered):
        if r != rank and recs:
            DIRECTORY.import_remote(recs)
