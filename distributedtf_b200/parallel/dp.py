"""Data-parallel member groups: one population member sharded over 2 or 4 GPUs.

Reference status: dormant — ``MirroredStrategy`` + ``AllReduceCrossTowerOps`` are wired up but
forced to one GPU (``resnet_run_loop.py:390-392``; ``distribution_utils.py:24-78``).  Here it is
live (BASELINE config 4): every replica runs forward/backward on its shard of the batch, then a
single kernel per replica (``dtf_allreduce_update``) reads every peer's gradient arena over
NVLink, sums, scales by 1/W, adds the regulariser gradient, applies the optimizer rule and
re-emits the bf16 GEMM weights.  A flag handshake in peer memory replaces both the NCCL
all-reduce and the host barrier:

    ready[r][p]  replica p's gradients of the current step are complete   (written into r's memory)
    done[r][p]   replica p has finished reading r's gradients             (gates r's next zeroing)

Both compare against the device-resident step counter, so the whole step replays from a CUDA
graph.  BatchNorm statistics stay local to each replica, as in the reference (nothing there
synchronises them across towers).  Per-device batch must divide the global batch
(``per_device_batch_size``, reference ``distribution_utils.py:58-78``).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from ..ops import native, native_ext
from . import symm

vp, ll, ull, i32, f32 = C.c_void_p, C.c_longlong, C.c_ulonglong, C.c_int, C.c_float
MAX_DP = 8
FLAG_BYTES = 1024          # ready[8] u64 | done[8] u64 | arrive counters | exchange counter | pad


class AllreduceUpdate(C.Structure):
    _fields_ = [("params", vp), ("slot0", vp), ("slot1", vp), ("pack", vp),
                ("peer_grads", vp * MAX_DP), ("peer_ready", vp * MAX_DP), ("peer_done", vp * MAX_DP),
                ("local_ready", vp), ("segs", vp), ("chunks", vp), ("hp", vp), ("step", vp),
                ("arrive_counter", vp), ("epoch_ctr", vp), ("epoch", ull), ("grad_scale", f32),
                ("world", i32), ("rank", i32), ("n_chunks", i32), ("timing", vp)]


native_ext.register("dtf_allreduce_update", [C.POINTER(AllreduceUpdate), vp])
native_ext.register("dtf_dp_step_begin", [vp, ll, vp, vp, vp, i32, ull, vp, vp])


def per_device_batch_size(batch_size: int, num_gpus: int) -> int:
    """Reference ``distribution_utils.per_device_batch_size``: the global batch must divide evenly."""
    if num_gpus <= 1:
        return batch_size
    remainder = batch_size % num_gpus
    if remainder:
        raise ValueError("When running with multiple GPUs, batch size must be a multiple of the number of "
                         "available GPUs. Found {} GPUs with a batch size of {}; try --batch_size={} instead."
                         .format(num_gpus, batch_size, batch_size - remainder))
    return batch_size // num_gpus


class DataParallelGroup:
    """Symmetric gradient arenas + flag blocks of the replicas of one member."""

    def __init__(self, grad_bytes: int, device: torch.device, group=None, local_peers: Optional[List["DataParallelGroup"]] = None):
        import torch.distributed as dist
        self.device = torch.device(device)
        self.group = group
        self.grad_buf = symm.DeviceBuffer(grad_bytes, self.device)
        self.flag_buf = symm.DeviceBuffer(FLAG_BYTES, self.device)
        if group is not None or (dist.is_available() and dist.is_initialized()):
            self.world = dist.get_world_size(group)
            self.rank = dist.get_rank(group)
            mine = (self.grad_buf.ipc_handle(), self.flag_buf.ipc_handle())
            allh: List = [None] * self.world
            dist.all_gather_object(allh, mine, group=group)
            self.peer_grads, self.peer_flags = [], []
            for r, (gh, fh) in enumerate(allh):
                if r == self.rank:
                    self.peer_grads.append(self.grad_buf.ptr)
                    self.peer_flags.append(self.flag_buf.ptr)
                else:
                    self.peer_grads.append(_open(gh))
                    self.peer_flags.append(_open(fh))
        else:
            self.world, self.rank = 1, 0
            self.peer_grads, self.peer_flags = [self.grad_buf.ptr], [self.flag_buf.ptr]

    def alloc_grads(self, nbytes: int) -> torch.Tensor:
        return self.grad_buf(nbytes)

    # flag block layout
    def ready_ptr(self, r: int) -> int:
        return self.peer_flags[r]

    def done_ptr(self, r: int) -> int:
        return self.peer_flags[r] + 8 * MAX_DP

    @property
    def arrive_update_ptr(self) -> int:
        return self.flag_buf.ptr + 16 * MAX_DP

    @property
    def arrive_begin_ptr(self) -> int:
        return self.flag_buf.ptr + 16 * MAX_DP + 64

    @property
    def timing_ptr(self) -> int:
        """Four u64 cells the all-reduce kernel accumulates its own device-side clock readings into."""
        return self.flag_buf.ptr + 16 * MAX_DP + 192

    def comm_stats(self) -> dict:
        """Device-timestamped accounting of the fused exchange since the group was built: time blocked on the
        peers' ready flags (the exposed communication), time of the remote gradient loads, kernel time."""
        return self.comm_stats_between(None, self.comm_raw())

    def comm_raw(self) -> list:
        """The four device counters of the exchange: ns blocked on ready flags, ns of remote loads, ns in the kernel,
        calls (accumulated by CTA 0 of every launch)."""
        return self.flag_buf.tensor[16 * MAX_DP + 192: 16 * MAX_DP + 224].view(torch.int64).tolist()

    @staticmethod
    def comm_stats_between(before, after) -> dict:
        cells = [a - (b if before is not None else 0) for a, b in zip(after, before or after)]
        calls = max(cells[3], 1)
        return {"calls": cells[3], "wait_us_per_call": cells[0] / calls / 1e3, "load_us_per_call": cells[1] / calls / 1e3,
                "kernel_us_per_call": cells[2] / calls / 1e3}

    @property
    def epoch_ctr_ptr(self) -> int:
        """Exchange counter of this replica: counts ``step_begin`` launches since the group was built.  It is
        deliberately not the member's step counter, which exploit and restore overwrite (members run different
        batch sizes, so their step counts differ) while ready[] / done[] keep the old epochs."""
        return self.flag_buf.ptr + 16 * MAX_DP + 128

    def step_begin(self, zero_region: torch.Tensor, step_dev: torch.Tensor) -> None:
        n = zero_region.numel() * zero_region.element_size() // 4
        native.check(native.lib().dtf_dp_step_begin(vp(zero_region.data_ptr()), n, vp(step_dev.data_ptr()),
                                                    vp(self.epoch_ctr_ptr), vp(self.done_ptr(self.rank)),
                                                    self.world, 0,
                                                    vp(self.arrive_begin_ptr), native.stream_ptr()),
                     "dtf_dp_step_begin")

    def allreduce_update(self, params, slot0, slot1, pack, segs_dev, chunks_dev, n_chunks, hp_dev, step_dev) -> None:
        p = AllreduceUpdate()
        p.params, p.slot0, p.slot1, p.pack = params.data_ptr(), slot0.data_ptr(), slot1.data_ptr(), pack.data_ptr()
        for r in range(self.world):
            p.peer_grads[r] = self.peer_grads[r]
            p.peer_ready[r] = self.ready_ptr(r)
            p.peer_done[r] = self.done_ptr(r)
        p.local_ready = self.ready_ptr(self.rank)
        p.segs, p.chunks, p.hp, p.step = segs_dev.data_ptr(), chunks_dev.data_ptr(), hp_dev.data_ptr(), step_dev.data_ptr()
        p.arrive_counter = self.arrive_update_ptr
        p.epoch_ctr = self.epoch_ctr_ptr
        p.epoch = 0
        p.grad_scale = 1.0 / self.world
        p.world, p.rank, p.n_chunks = self.world, self.rank, n_chunks
        p.timing = self.timing_ptr
        native.check(native.lib().dtf_allreduce_update(C.byref(p), native.stream_ptr()), "dtf_allreduce_update")


def _open(handle: bytes) -> int:
    out = vp()
    native.check(native.lib().dtf_ipc_open_handle(handle, C.byref(out)), "dtf_ipc_open_handle")
    return int(out.value)
