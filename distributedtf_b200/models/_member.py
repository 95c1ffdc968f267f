"""Shared machinery of the GPU-resident members (MNIST CNN, CIFAR-10 ResNet).

A member is *persistent*: it is built once (engine + resident dataset), keeps its
weights / optimizer slots / BN statistics on the device between rounds, and
writes its checkpoint at the end of every ``train`` call (for resume and for the
file-copy exploit baseline).  The reference instead rebuilds graph + session and
re-reads the checkpoint every round for every member
(``cifar10_main.py:321-330``; ``resnet_run_loop.py:397-398,452,463``).
"""
from __future__ import annotations

import math
import os
import time
from typing import Any, Dict, List, Optional

import torch

from .. import ckpt, report
from ..utils import nvtx
from ..model_base import ModelBase
from ..utils.data import BatchIterator, Dataset


def pick_device() -> torch.device:
    if not torch.cuda.is_available():
        return torch.device("cpu")
    idx = os.environ.get("DTF_LOCAL_GPU", os.environ.get("LOCAL_RANK", "0"))
    idx = int(idx) % torch.cuda.device_count()
    torch.cuda.set_device(idx)
    return torch.device("cuda", idx)


def resolve_engine(engine: str, device: torch.device) -> str:
    if engine != "auto":
        return engine
    if device.type == "cuda":
        from ..ops import native
        if native.available():
            return "native"
    return "eager"


class _TrainAccuracy:
    """Lazy ``train_accuracy`` of the last step: the head kernel counted the hits on the device
    (``correct_cell``); nothing is read back unless a hook actually logs the value."""

    def __init__(self, engine, batch: int):
        self._engine, self._batch = engine, max(int(batch), 1)

    def item(self) -> float:
        cell = getattr(self._engine, "correct_cell", None)
        if cell is None:
            acc = getattr(self._engine, "last_train_accuracy", None)
            return float(acc.item()) if acc is not None else float("nan")
        return float(cell.item()) / self._batch


class GpuMember(ModelBase):
    #: subclasses fill these in
    train_examples = 0
    eval_examples = 0
    csv_fields: List[str] = []

    def __init__(self, cluster_id: int, hparams: Dict[str, Any], save_base_dir: str, *,
                 engine: str = "auto", dtype: str = "bf16", synthetic: bool = True,
                 data_dir: str = "./datasets", max_train_steps: Optional[int] = None,
                 eval_batches: Optional[int] = None, resume: bool = False,
                 synthetic_mode: str = "random", device: Optional[torch.device] = None,
                 hooks: Optional[List[Any]] = None, dp_group=None, dp_rank: int = 0, dp_size: int = 1,
                 scoreboard: Optional[Dict[str, Any]] = None, benchmark=None, async_checkpoint: bool = False,
                 checkpoint_every: int = 1, epochs_between_evals: int = 1, stop_threshold: Optional[float] = None,
                 export_dir: Optional[str] = None, all_reduce_alg: Optional[str] = None, use_graphs: bool = True, **_):
        super().__init__(cluster_id, hparams, save_base_dir)
        self.device = device or pick_device()
        self.dp_group, self.dp_rank, self.dp_size = dp_group, int(dp_rank), int(dp_size)
        self._engine_requested = engine
        self.engine_kind = resolve_engine(engine, self.device)
        self.dtype = dtype
        self.synthetic = synthetic
        self.synthetic_mode = synthetic_mode
        self.data_dir = data_dir
        self.max_train_steps = max_train_steps
        self.eval_batches = eval_batches
        self.device_times: Dict[str, float] = {"train": 0.0, "eval": 0.0, "exploit": 0.0}
        self.hooks = list(hooks or [])          # utils.logs.hooks.Hook objects, called once per step
        self._bench_logger = None
        self.async_checkpoint = bool(async_checkpoint)   # native engines: save through the background writer
        self._ckpt_ticket = 0
        # 1 = every round (the reference's behaviour); k > 1 = every k-th round and the last; 0 = the last only.
        # The state lives on the GPU between rounds, so with a device exploit path files are only for resume.
        self.checkpoint_every = int(checkpoint_every)
        # run-loop controls of the reference harness (resnet_run_loop.py:434-448, 505-514)
        self.epochs_between_evals = max(1, int(epochs_between_evals))
        self.stop_threshold = stop_threshold
        self.export_dir = export_dir
        self.stopped_early = False
        self.all_reduce_alg = all_reduce_alg    # data-parallel member: fused_p2p (default) | nccl (library baseline)
        # native engines replay their step from a CUDA graph unless told otherwise (the launcher turns graphs off
        # for members that train concurrently inside one process, see launch.model_kwargs)
        self.use_graphs = bool(use_graphs)
        if benchmark and benchmark[0] == "BenchmarkFileLogger" and benchmark[1]:
            from ..utils.logs import logger as _logger
            self._bench_logger = _logger.BenchmarkFileLogger(os.path.join(benchmark[1], f"model_{cluster_id}"))
            self._bench_logger.log_run_info(type(self).__name__, "synthetic" if synthetic else "real",
                                            {"batch_size": int(hparams["batch_size"]), "dtype": dtype,
                                             "engine": engine, "synthetic_data": bool(synthetic)})
        self._round_batch_to_dp()
        self._board = scoreboard                # {"ptr": local pointer} or {"handle": CUDA IPC handle of the board}
        self._board_ptr: Optional[int] = None
        self.train_set, self.eval_set = self.build_datasets()
        self.symm = None                 # symmetric (peer-addressable) buffer, native engine on CUDA only
        self._pending_exploit: Optional[Dict[str, Any]] = None
        self.rounds_trained = 0
        self.engine = self.build_engine()
        self._attach_board()
        self._train_iter = BatchIterator(len(self.train_set), int(self.hparams["batch_size"]), True,
                                         seed=1000 + cluster_id, device=self.device)
        self.last_images_per_sec = 0.0
        if resume:
            self.restore()

    # ------------------------------------------------------------------ to implement
    def build_datasets(self):
        raise NotImplementedError

    def build_engine(self):
        raise NotImplementedError

    def preprocess(self, images_u8: torch.Tensor, training: bool) -> torch.Tensor:
        raise NotImplementedError

    def csv_row(self, epoch_index: int, accuracy: float) -> Dict[str, Any]:
        raise NotImplementedError

    def steps_per_epoch(self) -> int:
        return max(1, self.train_examples // int(self.hparams["batch_size"]))

    # ------------------------------------------------------------------ train / eval
    def _timed(self, key: str):
        member = self

        class _T:
            def __enter__(s):
                if member.device.type == "cuda":
                    s.e0 = torch.cuda.Event(enable_timing=True)
                    s.e1 = torch.cuda.Event(enable_timing=True)
                    s.e0.record()
                else:
                    s.t0 = time.time()
                return s

            def __exit__(s, *a):
                if member.device.type == "cuda":
                    s.e1.record()
                    s.e1.synchronize()
                    member.device_times[key] += s.e0.elapsed_time(s.e1) / 1e3
                else:
                    member.device_times[key] += time.time() - s.t0
        return _T()

    def _round_batch_to_dp(self) -> None:
        """A data-parallel member keeps its global batch a multiple of its replica count.  Sampled and explored
        batch sizes are odd about half the time; the reference's ``per_device_batch_size`` would reject those
        (``distribution_utils.py:58-78``) and the worker would then drop the member, so the value is rounded to the
        nearest lower multiple (upper at the bottom of the range) instead.  Every replica applies the same rule to
        the same hyper-parameters, so the group stays consistent."""
        if self.dp_size <= 1 or "batch_size" not in self.hparams:
            return
        bs = int(self.hparams["batch_size"])
        rounded = bs - bs % self.dp_size
        if rounded < 65:
            rounded += self.dp_size
        self.hparams["batch_size"] = rounded

    def train_one_epoch(self) -> int:
        bs = int(self.hparams["batch_size"])
        if self.dp_size > 1:          # reject before any step runs (reference distribution_utils.py:58-78)
            from ..parallel.dp import per_device_batch_size
            per_device_batch_size(bs, self.dp_size)
        self._train_iter.set_batch_size(bs)
        steps = self.steps_per_epoch()
        if self.max_train_steps is not None:
            steps = min(steps, self.max_train_steps)
        with nvtx.range(f"train epoch member {self.cluster_id}"), self._timed("train"):
            t0 = time.time()
            for _ in range(steps):
                idx = self._train_iter.next_indices()
                if self.dp_size > 1:          # same permutation on every replica; each takes its slice
                    per = bs // self.dp_size
                    idx = idx[self.dp_rank * per: (self.dp_rank + 1) * per]
                if self._indexed_steps():           # batch gathered inside the first kernel of the step
                    loss = self.engine.train_step_indexed(self.train_set.images, self.train_set.labels, idx)
                else:
                    x = self.preprocess(self.train_set.images[idx], True)
                    loss = self.engine.train_step(x, self.train_set.labels[idx])
                if self.hooks:                    # values stay on the device unless a hook logs them
                    self._global_step = getattr(self, "_global_step", 0) + 1
                    # the reference's default logged tensors (hooks_helper.py:33-35)
                    vals = {"cross_entropy": loss, "learning_rate": self.hparams["opt_case"]["lr"],
                            "train_accuracy": _TrainAccuracy(self.engine, int(idx.numel()))}
                    for h in self.hooks:
                        h.after_step(self._global_step, vals)
        # counted on replica 0 only, so that a data-parallel group reports its global batch once
        self.device_times["images"] = self.device_times.get("images", 0.0) + (steps * bs if self.dp_rank == 0 else 0)
        self.last_images_per_sec = steps * bs / max(time.time() - t0, 1e-9)
        return steps

    def _indexed_steps(self) -> bool:
        """Native engines read the batch out of the resident dataset by index (no gather / copy kernels)."""
        return (hasattr(self.engine, "train_step_indexed") and self.train_set.images.is_cuda
                and self.train_set.images.dtype == torch.uint8 and self.train_set.labels.dtype == torch.int32)

    def _board_pointer(self) -> Optional[int]:
        """Device address of the master's score board in this process (None: no board / not a CUDA member)."""
        if self._board is None or self.device.type != "cuda":
            return None
        if self._board_ptr is None:
            if "ptr" in self._board and int(self._board.get("owner_pid", os.getpid())) == os.getpid():
                self._board_ptr = int(self._board["ptr"])          # the board lives in this very process
            else:
                import ctypes as C
                from ..ops import native
                out = C.c_void_p()
                native.check(native.lib().dtf_ipc_open_handle(self._board["handle"], C.byref(out)), "open score board")
                self._board_ptr = int(out.value)
        return self._board_ptr

    def _attach_board(self) -> None:
        """Native single-GPU members: the evaluation kernels add their hits straight into the board slot.  Done
        once, before the first evaluation graph is captured (the pointer is baked into that graph)."""
        board = self._board_pointer()
        if board is not None and self.dp_size == 1 and hasattr(self.engine, "board_slot_ptr"):
            from ..parallel import scoreboard as sb
            self.engine.board_slot_ptr = sb.slot_ptr(board, self.cluster_id)

    def evaluate(self, final: bool = True) -> float:
        """Accuracy over the eval set (``final``: the last evaluation of this round -- only that one releases the
        round's epoch on the score board).  A data-parallel member shards the eval batches over its replicas
        (batch b goes to replica b % dp_size) and sums (correct, seen, non-finite-loss flag) over the group, so
        every replica ends with the same score and the same drop decision.

        With a score board the score is on the board (and the round's epoch released) by device-side work
        alone, enqueued before this function first synchronises with the host: the master ranks from the board
        while this member is still reading its own accuracy back (``pbt_cluster.py:114-128`` gathered it)."""
        from ..parallel import scoreboard as sb
        bs = int(self.hparams["batch_size"])
        it = BatchIterator(len(self.eval_set), bs, False, device=self.device)
        correct = torch.zeros((), dtype=torch.int64, device=self.device)
        seen = 0
        group = getattr(self.dp_group, "group", None) if self.dp_size > 1 else None
        shard = group is not None
        board = self._board_pointer() if self.dp_rank == 0 else None
        direct = board is not None and getattr(self.engine, "board_slot_ptr", 0) != 0
        epoch = self.rounds_trained + 1
        with nvtx.range(f"evaluate member {self.cluster_id}"), self._timed("eval"):
            if direct:
                sb.reset(board, self.cluster_id)
            for b, idx in enumerate(it.epoch_batches()):
                if self.eval_batches is not None and b >= self.eval_batches:
                    break
                if shard and b % self.dp_size != self.dp_rank:
                    continue
                if self._indexed_steps():
                    correct += self.engine.eval_batch_indexed(self.eval_set.images, self.eval_set.labels, idx).to(torch.int64)
                else:
                    x = self.preprocess(self.eval_set.images[idx], False)
                    correct += self.engine.eval_batch(x, self.eval_set.labels[idx]).to(torch.int64)
                seen += int(idx.numel())
            loss = self.engine.last_loss
            if direct and final:   # hits and seen are already in the slot; the drop decision is taken on the device
                sb.release(board, self.cluster_id, epoch, loss=loss.reshape(-1)[:1].float() if loss is not None
                           and loss.dtype == torch.float32 else None)
        bad = loss is not None and not bool(torch.isfinite(loss))
        if shard:
            import torch.distributed as dist
            packed = torch.stack([correct, torch.tensor(seen, device=self.device),
                                  torch.tensor(int(bad), device=self.device)]).to(torch.int64)
            dist.all_reduce(packed, group=group)
            correct, seen, bad = packed[0], int(packed[1].item()), bool(packed[2].item() > 0)
        if board is not None and bad and direct and not final:
            sb.release(board, self.cluster_id, epoch, dropped=True)
        if board is not None and not direct and (final or bad):
            if bad:
                sb.release(board, self.cluster_id, epoch, dropped=True)
            else:
                cell = correct.to(torch.int32).reshape(1).contiguous()
                sb.publish(board, self.cluster_id, cell, seen, epoch=epoch, reset=False)
        if bad:
            return float("nan")
        hits = int(correct.item())
        self._last_score = (hits, seen)
        return float(hits) / max(seen, 1)

    def publish_dropped(self) -> None:
        """The worker is about to drop this member (its training raised): leave the dropped marker on the
        score board so that the master's ranking does not wait for it."""
        if self.dp_rank != 0:
            return
        try:
            board = self._board_pointer()
            if board is not None:
                from ..parallel import scoreboard as sb
                sb.release(board, self.cluster_id, self.rounds_trained + 1, dropped=True)
        except Exception:
            pass        # a broken CUDA context cannot publish; the master falls back to the GET gather on timeout

    def _republish_score(self) -> None:
        """A member that stopped early (stop_threshold) keeps its score: show it for the new round as well."""
        board = self._board_pointer() if self.dp_rank == 0 else None
        if board is None:
            return
        from ..parallel import scoreboard as sb
        hits, seen = getattr(self, "_last_score", (0, 1))
        cell = torch.tensor([hits], dtype=torch.int32, device=self.device)
        sb.publish(board, self.cluster_id, cell, seen, epoch=self.rounds_trained + 1, reset=False)

    def make_symmetric_alloc(self, state_bytes: int) -> torch.Tensor:
        """``arena_alloc`` callback for native engines: puts the member state in exportable memory
        (state | snapshot 0 | snapshot 1 | control) and lists it in the symmetric directory."""
        from ..parallel import exploit, symm
        self.symm = exploit.MemberBuffer(state_bytes, self.device)
        symm.DIRECTORY.register_local(self.directory_key(self.cluster_id), self.symm.buf,
                                      int(os.environ.get("RANK", "0")))
        return self.symm.alloc_state(state_bytes)

    def directory_key(self, cluster_id: int) -> int:
        """Symmetric-directory key of replica ``dp_rank`` of member ``cluster_id`` (a loser replica pulls
        from the winner replica with the same data-parallel rank)."""
        return int(cluster_id) * 8 + self.dp_rank

    def train(self, num_epoch: int, total_epochs: int) -> None:
        """Train -> evaluate cycles of the reference's ``resnet_main`` (resnet_run_loop.py:446-507): every cycle
        trains ``epochs_between_evals`` epochs, evaluates, appends a CSV row, and stops early once the accuracy
        passes ``stop_threshold``.  Unlike the reference (``train_epochs // epochs_between_evals`` cycles, the
        remainder silently untrained) the last cycle also trains the remainder."""
        from ..utils.misc.model_helpers import past_stop_threshold
        self.flush_exploit()
        rows = []
        ebe = min(self.epochs_between_evals, max(num_epoch, 1))
        done = 0
        if self.stopped_early:
            self._republish_score()
        while done < num_epoch and not self.stopped_early:
            n = ebe if num_epoch - done - ebe >= ebe else num_epoch - done
            for _ in range(n):
                self.train_one_epoch()
            done += n
            if done >= num_epoch and self.symm is not None:
                # The round's weights are final: publish the snapshot BEFORE the evaluation, so that its epoch
                # flag is released ahead of the score's on the same stream -- by the time the master can rank
                # this member, a loser's pull of this snapshot never has to wait.
                from ..parallel import exploit
                exploit.publish(self, self.rounds_trained)
                self._published_round = self.rounds_trained
            self.accuracy = self.evaluate(final=done >= num_epoch)
            rows.append(self.csv_row(self.epoch_index_for_csv(done - 1), self.accuracy))
            if self._bench_logger is not None:      # resnet_run_loop.py:466 log_evaluation_result
                self._bench_logger.log_evaluation_result({"accuracy": self.accuracy, "global_step": self.engine.step})
            if not math.isnan(self.accuracy) and past_stop_threshold(self.stop_threshold, float(self.accuracy)):
                self.stopped_early = True           # resnet_run_loop.py:505-507; later train() calls are no-ops
                if done < num_epoch:
                    self._republish_score()         # the round ends here: release its epoch on the score board
        self.bump_epochs(num_epoch)
        if math.isnan(self.accuracy):
            return                      # the worker drops the member; nothing to persist
        if self.symm is not None and getattr(self, "_published_round", -1) != self.rounds_trained:
            from ..parallel import exploit      # (stopped early: the state did not change, the snapshot moves on)
            exploit.publish(self, self.rounds_trained)
        self.rounds_trained += 1
        last = self.epoches_trained >= total_epochs or self.stopped_early
        k = self.checkpoint_every
        if last or k == 1 or (k > 1 and self.rounds_trained % k == 0):
            self.save()
        report.append_csv(os.path.join(self.ensure_dir(), "learning_curve.csv"), self.csv_fields_now(), rows)
        if last and self.export_dir and self.dp_rank == 0:      # resnet_run_loop.py:510-514
            self.export(os.path.join(self.export_dir, "model_" + str(self.cluster_id)))

    export_model_name = "model"
    export_input_shape: List[int] = []

    def export_config(self) -> Dict[str, Any]:
        return {}

    def export(self, export_dir: str) -> str:
        """Inference bundle of this member (reference: ``classifier.export_savedmodel``,
        resnet_run_loop.py:510-514): parameters in the oracle's naming + ``signature.json``."""
        from ..utils.export import export as _export
        self.flush_exploit()
        eng = self.engine
        named = eng.export_named() if hasattr(eng, "export_named") else {k: v.detach() for k, v in eng.params.items()}
        return _export.export_inference_bundle(export_dir, named, self.export_model_name, self.export_config(),
                                               list(self.export_input_shape))

    def epoch_index_for_csv(self, e: int) -> int:
        return self.epoches_trained + e

    def bump_epochs(self, num_epoch: int) -> None:
        self.epoches_trained += num_epoch

    def csv_fields_now(self) -> List[str]:
        return list(self.csv_fields)

    # ------------------------------------------------------------------ state
    def on_hparams_changed(self) -> None:
        self._round_batch_to_dp()
        if self._pending_exploit is not None and not self._pending_exploit.get("explore_follows", False):
            self.flush_exploit()
        elif self._pending_exploit is None:
            self.engine.set_hparams(self.hparams)
        # else: SET arrived and EXPLORE follows; the fused kernel runs from perturb_hparams()

    def perturb_hparams(self) -> None:
        from .. import explore as _explore
        _explore.perturb_hparams(self.hparams, self._rng, tuple(self._perturb_factors))
        self._round_batch_to_dp()
        if self._pending_exploit is not None:
            self.flush_exploit()
        else:
            self.engine.set_hparams(self.hparams)

    def get_values(self):
        self.flush_exploit()
        return super().get_values()

    def flush_exploit(self) -> None:
        """Run a deferred device exploit: one kernel copies the winner's snapshot, writes this
        member's (possibly just explored) hyper-parameter block and resets its round cells."""
        extras, self._pending_exploit = self._pending_exploit, None
        if extras is None:
            return
        from ..parallel import exploit
        with nvtx.range(f"exploit member {self.cluster_id} <- {extras.get('src_id')}"), self._timed("exploit"):
            if extras.get("path") == "p2p" and hasattr(self.engine, "hparams_struct"):
                new_hp = self.engine.hparams_struct(self.hparams)
                exploit.receive(self, extras, new_hp=new_hp)
                self.engine.adopt_hparams(self.hparams)
            else:
                exploit.receive(self, extras)
                self.engine.set_hparams(self.hparams, keep_slots=True) if hasattr(self.engine, "adopt_hparams") \
                    else self.engine.set_hparams(self.hparams)

    def flush_io(self) -> None:
        """Barrier for a checkpoint handed to the background writer."""
        t, self._ckpt_ticket = self._ckpt_ticket, 0
        if t:
            ckpt.async_writer().wait(t)

    def save(self) -> None:
        nvtx.mark(f"checkpoint member {self.cluster_id}")
        self.flush_exploit()
        meta = {"model": type(self).__name__, "engine": self.engine_kind, "epoches_trained": self.epoches_trained}
        if self.async_checkpoint and self.device.type == "cuda" and hasattr(self.engine, "state"):
            self.flush_io()
            self._ckpt_ticket = ckpt.async_writer().save_state(
                self.ensure_dir(), self.engine.state, self.engine.state_tensors(), step=self.engine.step,
                optimizer=self.hparams["opt_case"]["optimizer"], meta=meta)
            return
        self.engine.synchronize()
        ckpt.save_checkpoint(self.ensure_dir(), self.engine.state_tensors(), step=self.engine.step,
                             optimizer=self.hparams["opt_case"]["optimizer"],
                             meta={"model": type(self).__name__, "engine": self.engine_kind,
                                   "epoches_trained": self.epoches_trained})

    def restore(self) -> bool:
        self.flush_io()
        loaded = ckpt.load_checkpoint(self.save_dir)
        if loaded is None:
            return False
        tensors, index = loaded
        self.engine.load_state(tensors, index["step"])
        self.epoches_trained = int(index.get("meta", {}).get("epoches_trained", self.epoches_trained))
        return True

    def adopt_state(self, extras: Dict[str, Any]) -> None:
        """Exploit: take over the winner's weights, slots, statistics and step."""
        if extras.get("src_id") == self.cluster_id:
            return
        path = extras.get("path", "file")
        if path == "file":
            with self._timed("exploit"):
                self.restore()          # the master already copied the winner's files here
        else:
            self._pending_exploit = dict(extras)     # runs fused with the explore write (flush_exploit)

    def donate_state(self, extras: Dict[str, Any]) -> None:
        from ..parallel import exploit
        with self._timed("exploit"):
            exploit.donate(self, extras)

    def drop(self) -> None:
        self.flush_io()
        super().drop()

    def close(self) -> None:
        self.flush_io()
        closer = getattr(self.engine, "close", None)
        if closer:
            closer()
