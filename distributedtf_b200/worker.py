"""The worker runtime: a blocking instruction loop that owns a slice of the population.

Parity: reference ``training_worker.py:12-105``.  Same handlers
(``add_graphs, train, get_all_values, set_values, explore_necessary_graphs``) and
the same member-drop rule: a member whose ``train`` raises, or whose accuracy is
NaN, is removed from the worker and its directory deleted; the master learns of
it through a shorter GET reply (``training_worker.py:63-80``).

B200-first differences: members are persistent GPU-resident objects (built once
in ``add_graphs``), SET records may carry ``extras`` describing a device-side
exploit (``[id, acc, hparams, extras]``), donor records (``extras["role"] ==
"src"``) ask a winner to take part in a two-sided transfer, and the profiling
reply may append a dict of device-timed phases.
"""
from __future__ import annotations

import math
import time
import traceback
from typing import Any, Callable, Dict, List, Optional

from .protocol import WorkerInstruction, validate


class TrainingWorker:
    def __init__(self, comm, master_rank: int, target_model_class: Callable[..., Any], *,
                 save_base_dir: str = "./savedata/model_", verbose: bool = True,
                 model_kwargs: Optional[Dict[str, Any]] = None,
                 fault_hook: Optional[Callable[[Any, int], None]] = None, mirror=None,
                 members_in_flight: int = 1, sm_limit: int = 0):
        self.worker_graphs: List[Any] = []
        self.is_expolore_only = False          # (sic) reference attribute name
        self.rank = comm.Get_rank()
        self.comm = comm
        self.master_rank = master_rank
        self.target_model_class = target_model_class
        self.save_base_dir = save_base_dir
        self.verbose = verbose
        self.model_kwargs = dict(model_kwargs or {})
        self.fault_hook = fault_hook           # test hook: f(member, round) may raise / poison
        self.mirror = mirror                   # DP group leader: callable(msg) that forwards to the followers
        # > 1: that many members of this worker train at once, each on its own CUDA stream and host thread,
        # (The reference trains them one after the other, training_worker.py:60-75.)
        self.members_in_flight = max(1, int(members_in_flight))
        self._member_streams: Dict[int, Any] = {}
        self.sm_limit = int(sm_limit)          # > 0: SMs each member's persistent kernels occupy (default: all)
        self.train_time = 0.0
        self.explore_time = 0.0
        self.exploit_time = 0.0
        self.rounds_seen = 0
        self.dropped: List[int] = []

    @property
    def is_explore_only(self) -> bool:
        return self.is_expolore_only

    def _say(self, *a) -> None:
        if self.verbose:
            print(*a, flush=True)

    # ------------------------------------------------------------------ loop
    def main_loop(self) -> None:
        while True:
            data = self.comm.recv(source=self.master_rank)
            inst = validate(data)
            if self.mirror is not None and inst in (WorkerInstruction.ADD_GRAPHS, WorkerInstruction.TRAIN,
                                                    WorkerInstruction.SET, WorkerInstruction.EXIT):
                self.mirror(data)
            if inst == WorkerInstruction.ADD_GRAPHS:
                self.is_expolore_only = data[3]
                self.add_graphs(data[1], data[2])
            elif inst == WorkerInstruction.TRAIN:
                self.train(data[1], data[2])
            elif inst == WorkerInstruction.GET:
                self.comm.send(self.get_all_values(), dest=self.master_rank)
            elif inst == WorkerInstruction.SET:
                self.set_values(data[1])
            elif inst == WorkerInstruction.EXPLORE:
                self.explore_necessary_graphs()
            elif inst == WorkerInstruction.GET_PROFILING_INFO:
                self.comm.send(self.profiling_info(), dest=self.master_rank)
            elif inst == WorkerInstruction.EXIT:
                self.shutdown()
                break

    # ------------------------------------------------------------------ handlers
    def add_graphs(self, hparam_list: List[dict], id_begin: int) -> None:
        self._say("[{}]Got {} hparams".format(self.rank, len(hparam_list)))
        if self.sm_limit > 0:       # baked into the members' CUDA graphs at capture
            try:
                from .ops import native
                native.set_sm_limit(self.sm_limit)
            except Exception:
                pass
        for offset, hparam in enumerate(hparam_list):
            member = self.target_model_class(id_begin + offset, hparam, self.save_base_dir,
                                             **self.model_kwargs)
            self.worker_graphs.append(member)
        self._exchange_symmetric_directory()

    def _exchange_symmetric_directory(self) -> None:
        """Publish the CUDA IPC handles of this worker's members to the other workers (device
        exploit paths).  A collective over the worker process group; skipped without one."""
        bufs = {(g.directory_key(g.cluster_id) if hasattr(g, "directory_key") else g.cluster_id): g.symm.buf
                for g in self.worker_graphs if getattr(g, "symm", None) is not None}
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                from .parallel import symm
                symm.exchange_directory(bufs)
        except ImportError:
            pass

    def _train_member(self, g, num_epoches: int, total_epochs: int) -> bool:
        """Trains one member; returns False when it has to be dropped (exception or NaN score)."""
        try:
            if self.fault_hook is not None:
                self.fault_hook(g, self.rounds_seen)
            g.train(num_epoches, total_epochs)
            self._say("Model {} epoch = {},  acc = {}".format(
                g.cluster_id, g.epoches_trained, g.get_accuracy()))
            acc = g.get_accuracy()
            return not (isinstance(acc, float) and math.isnan(acc))
        except Exception:  # member-level failure: drop it, keep the worker alive
            traceback.print_exc()          # always: a dropped member must leave a trace (the reference prints one too)
            marker = getattr(g, "publish_dropped", None)
            if marker is not None:      # device score board: the master must not wait for this member
                marker()
            return False

    def _train_member_on_stream(self, g, num_epoches: int, total_epochs: int) -> bool:
        import torch
        dev = getattr(g, "device", None)
        if dev is None or getattr(dev, "type", "cpu") != "cuda":
            return self._train_member(g, num_epoches, total_epochs)
        try:
            torch.cuda.set_device(dev)
            st = self._member_streams.get(g.cluster_id)
            if st is None:
                from .utils.streams import new_stream
                st = self._member_streams[g.cluster_id] = new_stream(dev)
            st.wait_stream(torch.cuda.default_stream(dev))       # exploit / explore writes of the main thread
            with torch.cuda.stream(st):
                ok = self._train_member(g, num_epoches, total_epochs)
            st.synchronize()
            return ok
        except Exception:
            if self.verbose:
                traceback.print_exc()
            return False

    def train(self, num_epoches: int, total_epochs: int) -> None:
        t0 = time.time()
        k = min(self.members_in_flight, len(self.worker_graphs))
        if k > 1:
            from concurrent.futures import ThreadPoolExecutor
            with ThreadPoolExecutor(max_workers=k, thread_name_prefix="dtf-member") as pool:
                oks = list(pool.map(lambda g: self._train_member_on_stream(g, num_epoches, total_epochs),
                                    list(self.worker_graphs)))
        else:
            oks = [self._train_member(g, num_epoches, total_epochs) for g in list(self.worker_graphs)]
        failed = [g for g, ok in zip(list(self.worker_graphs), oks) if not ok]
        for g in failed:
            self._say("Error occured , graph {} removed".format(g.cluster_id))
            self.dropped.append(g.cluster_id)
            try:
                g.drop()
            finally:
                self.worker_graphs.remove(g)
        self.rounds_seen += 1
        self.train_time += time.time() - t0
        if self.mirror is not None:            # the followers drop exactly what the leader dropped
            self.mirror(("DROP", [g.cluster_id for g in failed]))

    def get_all_values(self) -> List[List[Any]]:
        return [g.get_values() for g in self.worker_graphs]

    def set_values(self, values_to_set: List[List[Any]]) -> None:
        t0 = time.time()
        for v in values_to_set:
            extras = v[3] if len(v) > 3 else None
            for g in self.worker_graphs:
                if g.cluster_id != v[0]:
                    continue
                if extras and extras.get("role") == "src":
                    donate = getattr(g, "donate_state", None)
                    if donate is not None:
                        donate(extras)
                else:
                    g.set_values(v)
                    g.need_explore = True
        self.exploit_time += time.time() - t0

    def explore_necessary_graphs(self) -> None:
        t0 = time.time()
        explored: Dict[int, Any] = {}
        for g in self.worker_graphs:
            if g.need_explore or self.is_expolore_only:
                self._say("[{}]Exploring graph {}".format(self.rank, g.cluster_id))
                g.perturb_hparams()
                g.need_explore = False
                explored[g.cluster_id] = g.hparams
        if self.mirror is not None:            # followers adopt the leader's draw instead of drawing their own
            self.mirror(("HPARAMS", explored))
        self.explore_time += time.time() - t0

    def profiling_info(self) -> List[Any]:
        device: Dict[str, float] = {}
        for g in self.worker_graphs:
            for k, val in getattr(g, "device_times", {}).items():
                device[k] = device.get(k, 0.0) + float(val)
        device["exploit_worker"] = self.exploit_time
        return [self.train_time, self.explore_time, device]

    def shutdown(self) -> None:
        for g in self.worker_graphs:
            close = getattr(g, "close", None)
            if close is not None:
                close()


class FollowerWorker(TrainingWorker):
    """A non-leader replica of a data-parallel member group.  It never talks to the master: the leader
    mirrors ADD_GRAPHS / TRAIN / SET / EXIT plus two internal records — ("HPARAMS", {id: hparams}) after an
    explore so every replica holds the same draw, and ("DROP", [ids]) after a round."""

    def __init__(self, recv_from_leader: Callable[[], Any], target_model_class, **kw):
        class _NoComm:
            def Get_rank(self):
                return -1
        super().__init__(_NoComm(), 0, target_model_class, **kw)
        self._recv = recv_from_leader

    def main_loop(self) -> None:
        import copy
        while True:
            data = self._recv()
            tag = data[0]
            if tag == WorkerInstruction.ADD_GRAPHS:
                self.is_expolore_only = data[3]
                self.add_graphs(data[1], data[2])
            elif tag == WorkerInstruction.TRAIN:
                for g in self.worker_graphs:
                    try:
                        g.train(data[1], data[2])
                    except Exception:
                        if self.verbose:
                            traceback.print_exc()
            elif tag == WorkerInstruction.SET:
                self.set_values(data[1])
            elif tag == "HPARAMS":
                for g in self.worker_graphs:
                    if g.cluster_id in data[1]:
                        g.hparams = copy.deepcopy(data[1][g.cluster_id])
                        g.need_explore = False
                        if getattr(g, "_pending_exploit", None) is not None:
                            g.flush_exploit()
                        else:
                            g.on_hparams_changed()
            elif tag == "DROP":
                for g in [g for g in self.worker_graphs if g.cluster_id in data[1]]:
                    self.worker_graphs.remove(g)
            elif tag == WorkerInstruction.EXIT:
                self.shutdown()
                break
