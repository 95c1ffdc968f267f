"""Machinery shared by the native (sm_100a) engines: the flat state arena, the device-resident
hyper-parameter block, CUDA-graph capture per batch size, the step API and the state API.

Arena layout::

    [ hparams | step | params fp32 | slot0 | slot1 | running statistics ]   <- exploit copies this
    [ bf16 GEMM-ready weights ]                                            <- derived, refreshed by the optimizer
    [ grads fp32 | per-step statistics | loss, correct ]                    <- zeroed every step

A subclass describes its parameters as ``native.OptSegment`` records (``self.segments``), allocates
its activations, and implements ``_train_kernels`` / ``_eval_kernels`` as a sequence of native launches.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Any, Dict, List, Optional, Tuple

import torch

from ..ops import native, reference
from ..utils.streams import new_stream


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


from ..utils.capture_gate import GATE as _GATE      # steps run concurrently, a CUDA-graph capture runs alone


class NativeEngineBase:
    kind = "native"
    image_channels = 3
    image_size = 32

    def __init__(self, device, *, max_batch: int, use_graphs: bool, arena_alloc=None, dp_group=None, seed: int = 0,
                 all_reduce_alg: Optional[str] = None):
        native.require()
        self.device = torch.device(device)
        self.max_batch = max_batch
        self.use_graphs = use_graphs and os.environ.get("DTF_NO_GRAPHS") != "1"
        self.hp: Dict[str, Any] = {}
        self._graphs: Dict[Tuple[str, int], torch.cuda.CUDAGraph] = {}
        self.dp_group = dp_group          # torch.distributed group of this member's replicas (None = 1 GPU)
        self.dp = None
        # gradient exchange of a data-parallel member: "fused_p2p" = the one-shot all-reduce + update kernel over
        # peer memory (the product); "nccl" = library all-reduce followed by the plain optimizer kernel (the
        # baseline the reference's dormant MirroredStrategy path amounts to, distribution_utils.py:41-47)
        self.all_reduce_alg = (all_reduce_alg or "fused_p2p") if dp_group is not None else None
        if self.all_reduce_alg not in (None, "fused_p2p", "nccl"):
            raise ValueError("all_reduce_alg of the native engine must be fused_p2p or nccl")
        self._arena_alloc = arena_alloc
        self.last_loss: Optional[torch.Tensor] = None
        self.launches_per_step = 0
        self._pp_counter = 0
        self._seed = seed
        self._count = 0
        self._side = None                       # side stream for independent kernels (engine-specific use)
        self._cap_stream = None                 # stream CUDA graphs are captured on
        self.board_slot_ptr = 0                 # this member's score-board slot (device / peer pointer), 0 = none
        # static loss scale (reference resnet_run_loop.py:284-294): the head kernel multiplies d(loss)/d(logits) by it,
        # the optimizer / all-reduce kernel multiplies the gradients by 1 / scale (DtfHparams.inv_loss_scale)
        self.loss_scale = 1.0
        self._src = None                        # (images_all, labels_all) while an indexed step is being issued
        self._idx_buf = None

    # ------------------------------------------------------------------ memory
    def _finish_segments(self) -> None:
        chunks: List[int] = []
        for si, s in enumerate(self.segments):
            for start in range(0, s.count, native.OPT_CHUNK):
                chunks += [si, start]
        self.n_chunks = len(chunks) // 2
        self._chunks_host = torch.tensor(chunks, dtype=torch.int32)

    def _alloc_arena(self) -> None:
        dev = self.device
        P = _align(self.n_params * 4) // 4
        self.P_pad = P
        HP = _align(C.sizeof(native.Hparams))
        off = {}
        cur = 0
        for name, nbytes in [("hparams", HP), ("step", 256), ("params", P * 4), ("slot0", P * 4), ("slot1", P * 4),
                             ("running", _align(max(self.n_running, 1) * 4))]:
            off[name] = cur
            cur += _align(nbytes)
        self.state_bytes_total = cur
        alloc = self._arena_alloc or (lambda n: torch.zeros(n, dtype=torch.uint8, device=dev))
        self.state = alloc(cur)[:cur]                 # the region exploit copies
        self.hp_state_offset = off["hparams"]
        self._state_off = off
        v = self.state

        def view(name, dtype, count):
            esz = torch.empty(0, dtype=dtype).element_size()
            return v[off[name]: off[name] + count * esz].view(dtype)

        self.hp_dev = v[off["hparams"]: off["hparams"] + C.sizeof(native.Hparams)]
        self.step_dev = view("step", torch.int64, 1)
        self.params = view("params", torch.float32, P)
        self.slot0 = view("slot0", torch.float32, P)
        self.slot1 = view("slot1", torch.float32, P)
        self.running = view("running", torch.float32, self.n_running)
        # GEMM-ready copies of the conv kernels in the operand type of the engine (bf16, or fp32 for tf32 math)
        self.pack = torch.zeros(max(self.n_pack, 8), dtype=getattr(self, "act_dtype", torch.bfloat16), device=dev)
        # per-step zeroed region: grads | batch statistics | loss, correct
        zfloats = P + _align(max(self.n_stats, 1) * 4) // 4 + 64
        if self.dp_group is not None and self.all_reduce_alg == "fused_p2p":
            # data-parallel member: the gradient arena must be addressable by the peer replicas
            from ..parallel.dp import DataParallelGroup
            self.dp = DataParallelGroup(zfloats * 4, dev, group=self.dp_group)
            self.zero_region = self.dp.alloc_grads(zfloats * 4).view(torch.float32)
        else:
            self.zero_region = torch.zeros(zfloats, dtype=torch.float32, device=dev)
        self.grads = self.zero_region[:P]
        s0 = P
        self.stats = self.zero_region[s0: s0 + self.n_stats]
        m0 = s0 + _align(max(self.n_stats, 1) * 4) // 4
        self.loss_cell = self.zero_region[m0: m0 + 1]
        self.correct_cell = self.zero_region[m0 + 1: m0 + 2].view(torch.int32)
        # evaluation accumulators live outside the zeroed region; exploit clears them for the loser
        self.eval_cells = torch.zeros(4, dtype=torch.float32, device=dev)
        self.round_cells = self.eval_cells
        self.segs_dev = native.struct_to_tensor((native.OptSegment * len(self.segments))(*self.segments), dev)
        self.chunks_dev = self._chunks_host.to(dev)
        self.in_u8 = torch.zeros(self.max_batch, self.image_size, self.image_size, self.image_channels,
                                 dtype=torch.uint8, device=dev)
        self.labels = torch.zeros(self.max_batch, dtype=torch.int32, device=dev)

    def repack(self) -> None:
        native.pack_weights(self.params, self.pack, self.segs_dev, self.chunks_dev, self.n_chunks)

    # ------------------------------------------------------------------ hparams
    def lr_schedule(self, hp: Dict[str, Any]):
        """(boundaries, values) of the piecewise-constant schedule; constant by default."""
        lr = float(hp["opt_case"]["lr"])
        return [], [lr]

    def hparams_struct(self, hp: Dict[str, Any]) -> "native.Hparams":
        oc = hp["opt_case"]
        h = native.Hparams()
        h.optimizer = reference.OPT_IDS[oc["optimizer"]]
        h.regularizer = reference.REG_IDS[hp.get("regularizer", "None")] if self.uses_regularizer else 0
        h.momentum = float(oc.get("momentum", 0.0))
        h.grad_decay = float(oc.get("grad_decay", 0.9))
        h.weight_decay = float(hp.get("weight_decay", 0.0))
        h.inv_loss_scale = 1.0 / float(getattr(self, "loss_scale", 1.0) or 1.0)
        bounds, values = self.lr_schedule(hp)
        if len(bounds) > native.MAX_LR_BOUNDARIES:
            raise ValueError("LR schedule has too many boundaries for the device block")
        h.n_boundaries = len(bounds)
        h.batch_size = int(hp["batch_size"])
        for i, b in enumerate(bounds):
            h.boundaries[i] = int(b)
        for i, v in enumerate(values):
            h.lr_values[i] = float(v)
        return h

    uses_regularizer = True

    def adopt_hparams(self, hp: Dict[str, Any]) -> None:
        """The device block was already written by the exploit kernel: only mirror it on the host."""
        self.hp = hp
        self._hp_struct = self.hparams_struct(hp)

    def set_hparams(self, hp: Dict[str, Any], reset_slots: bool = False, keep_slots: bool = False) -> None:
        prev = self.hp.get("opt_case", {}).get("optimizer") if self.hp else None
        self.hp = hp
        oc = hp["opt_case"]
        h = self.hparams_struct(hp)
        self._hp_struct = h
        self.hp_dev.copy_(native.struct_to_tensor(h, "cpu").to(self.device, non_blocking=True))
        if reset_slots or (prev != oc["optimizer"] and not keep_slots):
            a, b = reference.slot_init_values(oc["optimizer"])
            native.fill(self.slot0, a)
            native.fill(self.slot1, b)

    @property
    def step(self) -> int:
        return int(self.step_dev.item())

    # ------------------------------------------------------------------ step plumbing
    def _begin_step(self) -> None:
        if self.dp is not None:
            self.dp.step_begin(self.zero_region, self.step_dev)
        else:
            native.step_begin(self.zero_region, self.step_dev, bump=True)
        self._count += 1

    def _nccl_world(self) -> int:
        import torch.distributed as dist
        return dist.get_world_size(self.dp_group)

    def _apply_update(self) -> None:
        if self.all_reduce_alg == "nccl":
            # library baseline: NCCL all-reduce of the gradient arena, then the single-GPU optimizer kernel with
            # a 1/W scale.  Issued outside the step's CUDA graph (see _run).
            if torch.cuda.is_current_stream_capturing():
                self.launches_per_step = self._count + 1
                return
            import torch.distributed as dist
            dist.all_reduce(self.grads, group=self.dp_group)
            native.optimizer_step(self.params, self.grads, self.slot0, self.slot1, self.pack, self.segs_dev,
                                  self.chunks_dev, self.n_chunks, self.hp_dev, self.step_dev, 1.0 / self._nccl_world())
            self._count += 1
            self.launches_per_step = self._count
            return
        if self.dp is not None:
            self.dp.allreduce_update(self.params, self.slot0, self.slot1, self.pack, self.segs_dev,
                                     self.chunks_dev, self.n_chunks, self.hp_dev, self.step_dev)
        else:
            native.optimizer_step(self.params, self.grads, self.slot0, self.slot1, self.pack, self.segs_dev,
                                  self.chunks_dev, self.n_chunks, self.hp_dev, self.step_dev, 1.0)
        self._count += 1
        self.launches_per_step = self._count

    def _train_kernels(self, nb: int, pp_mode: int) -> None:
        raise NotImplementedError

    def _eval_kernels(self, nb: int) -> None:
        raise NotImplementedError

    def _run(self, kind: str, nb: int, fn) -> None:
        if not self.use_graphs:
            fn()
            return
        key = (kind, nb)
        g = self._graphs.get(key)
        if g is None:
            # One capture at a time per process, with no other member's step in flight, in thread-local error
            # mode (other members may still be evaluating / saving on their own streams and threads).
            _GATE.enter_capture()
            try:
                fn()                               # first use: a plain run (sets kernel attributes) ...
                cur = torch.cuda.current_stream(self.device)
                cur.synchronize()
                # ... then capture, on a side stream, with the graph object's own begin / end calls: the
                # torch.cuda.graph context manager would add a device-wide synchronize, gc.collect() and
                # empty_cache() to every capture (~100 ms in a PBT worker, paid again whenever explore changes a
                # member's batch size)
                g = torch.cuda.CUDAGraph()
                if self._cap_stream is None:
                    self._cap_stream = new_stream(self.device)
                self._cap_stream.wait_stream(cur)
                with torch.cuda.stream(self._cap_stream):
                    g.capture_begin(capture_error_mode="thread_local")
                    try:
                        fn()
                    finally:
                        g.capture_end()
                cur.wait_stream(self._cap_stream)
            finally:
                _GATE.exit_capture()
            self._graphs[key] = g
            if len(self._graphs) > 16:
                self._graphs.pop(next(iter(self._graphs)))
            return                                 # the plain run already performed this step
        _GATE.enter_step()
        try:
            g.replay()
        finally:
            _GATE.exit_step()
        name = kind[0] if isinstance(kind, tuple) else kind
        if self.all_reduce_alg == "nccl" and str(name).startswith("train"):
            self._apply_update()                   # the captured graph stops before the library all-reduce

    def load_batch(self, images_u8: torch.Tensor, labels: torch.Tensor) -> int:
        nb = int(images_u8.shape[0])
        if nb > self.max_batch:
            raise ValueError(f"batch {nb} exceeds the engine's max_batch {self.max_batch}")
        self.in_u8[:nb].copy_(images_u8, non_blocking=True)
        self.labels[:nb].copy_(labels.to(torch.int32), non_blocking=True)
        return nb

    # -- steps that read their batch straight out of a resident dataset: the preprocessing kernel gathers images
    #    and labels by index, so a step launches no torch gather / copy kernels at all
    def _load_indices(self, images_all: torch.Tensor, labels_all: torch.Tensor, idx: torch.Tensor) -> int:
        nb = int(idx.numel())
        if nb > self.max_batch:
            raise ValueError(f"batch {nb} exceeds the engine's max_batch {self.max_batch}")
        if labels_all.dtype != torch.int32 or images_all.dtype != torch.uint8:
            raise TypeError("indexed steps need a uint8 image tensor and int32 labels")
        if self._idx_buf is None:
            self._idx_buf = torch.zeros(self.max_batch, dtype=torch.int32, device=self.device)
        self._idx_buf[:nb].copy_(idx, non_blocking=True)
        self._src = (images_all, labels_all)
        return nb

    def preprocess_batch(self, nb: int, mode: int, counter: int = 0) -> None:
        """First kernel of every step: uint8 batch (copied in, or gathered by index) -> standardised bf16 image."""
        if self._src is not None:
            images_all, labels_all = self._src
            native.preprocess(images_all, self._idx_buf, self.image, mode=mode, seed=self._seed, counter=counter, n=nb,
                              labels_all=labels_all, labels_out=self.labels, step=self.step_dev)
        else:
            native.preprocess(self.in_u8, None, self.image, mode=mode, seed=self._seed, counter=counter, n=nb,
                              step=self.step_dev)

    def train_step_indexed(self, images_all: torch.Tensor, labels_all: torch.Tensor, idx: torch.Tensor,
                           augment: bool = True) -> torch.Tensor:
        nb = self._load_indices(images_all, labels_all, idx)
        mode = self.train_pp_mode if augment else self.eval_pp_mode
        self._pp_counter += nb
        try:
            self._run(("train_idx" if augment else "train_idx_noaug", images_all.data_ptr(), labels_all.data_ptr()), nb,
                      lambda: self._train_kernels(nb, mode))
        finally:
            self._src = None
        self.last_loss = self.loss_cell
        return self.loss_cell

    def eval_batch_indexed(self, images_all: torch.Tensor, labels_all: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        nb = self._load_indices(images_all, labels_all, idx)
        before = self.eval_cells[1:2].view(torch.int32).clone()
        try:
            self._run(("eval_idx", images_all.data_ptr(), labels_all.data_ptr()), nb, lambda: self._eval_kernels(nb))
        finally:
            self._src = None
        return (self.eval_cells[1:2].view(torch.int32) - before).sum()

    def train_step_u8(self, images_u8: torch.Tensor, labels: torch.Tensor, augment: bool = True) -> torch.Tensor:
        """uint8 NHWC images (device or pinned host) -> one optimizer step; returns the loss cell."""
        nb = self.load_batch(images_u8, labels)
        mode = self.train_pp_mode if augment else self.eval_pp_mode
        self._pp_counter += nb
        self._run("train" if augment else "train_noaug", nb, lambda: self._train_kernels(nb, mode))
        self.last_loss = self.loss_cell
        return self.loss_cell

    train_pp_mode, eval_pp_mode = 2, 1

    def eval_batch_u8(self, images_u8: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        nb = self.load_batch(images_u8, labels)
        before = self.eval_cells[1:2].view(torch.int32).clone()
        self._run("eval", nb, lambda: self._eval_kernels(nb))
        return self.eval_cells[1:2].view(torch.int32) - before

    # GpuMember hooks: preprocessing happens inside the step, so these take raw uint8 batches
    def preprocess(self, images_u8: torch.Tensor, training: bool) -> torch.Tensor:
        return images_u8

    def train_step(self, images_u8: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        return self.train_step_u8(images_u8, labels, augment=True)

    def eval_batch(self, images_u8: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        return self.eval_batch_u8(images_u8, labels).sum()

    # ------------------------------------------------------------------ state
    def state_tensors(self) -> Dict[str, torch.Tensor]:
        return {"params": self.params[: self.n_params], "slot0": self.slot0[: self.n_params],
                "slot1": self.slot1[: self.n_params], "running": self.running, "step": self.step_dev}

    @torch.no_grad()
    def load_state(self, tensors: Dict[str, Any], step: int) -> None:
        for k, dst in self.state_tensors().items():
            if k == "step":
                continue
            src = tensors[k]
            src = src if isinstance(src, torch.Tensor) else torch.from_numpy(src)
            dst.copy_(src.to(self.device).view_as(dst))
        self.step_dev.fill_(int(step))
        self.repack()

    def state_bytes(self) -> int:
        return self.state_bytes_total

    def after_state_replaced(self) -> None:
        """Call after the state arena was overwritten from outside (exploit): refresh derived data."""
        self.repack()

    def synchronize(self) -> None:
        # stream-scoped on purpose: a device-wide synchronize is illegal while another member's thread is
        # capturing a CUDA graph (worker.members_in_flight) and would invalidate that capture
        torch.cuda.current_stream(self.device).synchronize()
        if self._side is not None:
            self._side.synchronize()

    def close(self) -> None:
        self._graphs.clear()
        from ..utils import streams
        for name in ("_side", "_cap_stream"):
            st = getattr(self, name, None)
            if st is not None:
                st.synchronize()
                streams.release(st)
                setattr(self, name, None)
