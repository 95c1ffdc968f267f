"""The PBT controller (runs on the master endpoint only).

Parity: reference ``pbt_cluster.py:27-266`` — population bookkeeping, the round
loop (TRAIN fan-out -> exploit -> explore), truncation-selection exploit
(bottom ceil(pop/4) <- top ceil(pop/4), pairwise), gathers, JSON dumps and the
profiling summary.  Method names follow the reference so driver scripts port
over unchanged.

What is different, on purpose:

* **Exploit transport is pluggable** (``exploit_path``): ``"file"`` reproduces
  the reference (checkpoint files copied by the master, ``pbt_cluster.py:168-181``);
  ``"p2p"`` makes the loser's GPU pull the winner's published snapshot over
  NVLink peer memory with a hand-written kernel; ``"nccl"`` and ``"host"`` are
  the library baselines.  For the device paths the master only names the pair in
  the SET record (``[id, acc, hparams, extras]``); no tensor ever crosses the
  control plane.
* SET goes only to workers that own a loser (or, for two-sided paths, a donor) —
  the reference also isends to the master's own rank, a message nobody receives
  (``pbt_cluster.py:154-162``).
* ``recv`` carries an optional timeout so a dead worker aborts the run instead of
  hanging it.
* With a device score board (``parallel/scoreboard.py``) the board IS the round barrier and
  the source of the scores: after TRAIN the master polls it until every live member's evaluation
  kernels have released the round's epoch, ranks, and sends SET — without a GET round trip on
  that path (the reference's GET is both gather and barrier, ``pbt_cluster.py:114-128,191-193``).
  The hyper-parameter dicts still travel by GET, but off the critical path: EXPLORE, GET and the
  next round's TRAIN are sent back to back, so a worker answers and starts training at once.
"""
from __future__ import annotations

import copy
import datetime
import math
import os
import random
import time
from typing import Any, Callable, Dict, List, Optional, Sequence

from . import ckpt, protocol, report
from .utils import nvtx
from .hpspace import generate_random_hparam
from .protocol import WorkerInstruction

EXPLOIT_PATHS = ("file", "p2p", "nccl", "host")
TWO_SIDED_PATHS = ("nccl", "host")


def partition(pop_size: int, n_workers: int) -> List[range]:
    """Contiguous member slices per worker: ceil(pop / workers) each, last ones
    possibly short or empty (reference ``pbt_cluster.py:56-75``)."""
    if n_workers < 1:
        raise ValueError("need at least one worker")
    per = int(math.ceil(float(pop_size) / float(n_workers)))
    out = []
    for k in range(n_workers):
        begin = min(k * per, pop_size)
        out.append(range(begin, min(begin + per, pop_size)))
    return out


def exploit_pairs(sorted_ids: Sequence[Any]) -> List[tuple]:
    """(winner_pos, loser_pos) index pairs into an ascending-by-score list."""
    n = int(math.ceil(len(sorted_ids) / 4.0))
    total = len(sorted_ids)
    return [(total - n + i, i) for i in range(n)]


class PBTCluster:
    def __init__(self, pop_size: int, comm, master_rank: int, epochs_per_round: int,
                 do_exploit: bool = True, do_explore: bool = True, *,
                 save_root: str = "./savedata", exploit_path: str = "file",
                 rng: Optional[random.Random] = None,
                 recv_timeout: Optional[float] = None,
                 initial_hparams: Optional[List[Dict[str, Any]]] = None,
                 scoreboard=None, verbose: bool = True,
                 hparam_factory: Callable[..., Dict[str, Any]] = generate_random_hparam):
        if exploit_path not in EXPLOIT_PATHS:
            raise ValueError(f"exploit_path must be one of {EXPLOIT_PATHS}")
        self.pop_size = pop_size
        self.comm = comm
        self.master_rank = master_rank
        self.epochs_per_round = epochs_per_round
        self.do_exploit = do_exploit
        self.do_explore = do_explore
        self.save_root = save_root
        self.exploit_path = exploit_path
        self.recv_timeout = recv_timeout
        self.scoreboard = scoreboard
        self.verbose = verbose
        self._rng = rng
        self._hparam_factory = hparam_factory

        self.exploit_time = 0.0          # rank + transfer + SET (not the gather wait)
        self.gather_time = 0.0           # time blocked in GET replies (slowest worker)
        self.exploit_bytes = 0
        self.rounds_done = 0
        self.history: List[Dict[str, Any]] = []   # per-round pairs, for tests / reports
        self.cluster_id_to_worker_rank: Dict[int, int] = {}
        self._known: Optional[Dict[int, List[Any]]] = None   # last GET view: id -> [id, acc, hparams]
        self.board_timeout = 120.0
        self.board_fallbacks = 0                  # rounds ranked from a GET gather because the board timed out

        self.dispatch_hparams_to_workers(initial_hparams)

    # ------------------------------------------------------------------ helpers
    def _say(self, *a) -> None:
        if self.verbose:
            print(*a, flush=True)

    @property
    def worker_ranks(self) -> List[int]:
        return [i for i in range(self.comm.Get_size()) if i != self.master_rank]

    def _broadcast(self, msg: tuple) -> None:
        reqs = [self.comm.isend(msg, dest=i) for i in self.worker_ranks]
        for r in reqs:
            r.wait()

    def _gather(self) -> Dict[int, Any]:
        out = {}
        for i in self.worker_ranks:
            out[i] = self.comm.recv(source=i, timeout=self.recv_timeout) \
                if self.recv_timeout is not None else self.comm.recv(source=i)
        return out

    # ------------------------------------------------------------------ set-up
    def dispatch_hparams_to_workers(self, initial_hparams=None) -> None:
        if initial_hparams is not None:
            hps = [copy.deepcopy(h) for h in initial_hparams]
            self.pop_size = len(hps)
        else:
            hps = [self._hparam_factory(self._rng) if self._rng is not None
                   else self._hparam_factory() for _ in range(self.pop_size)]
        self._say("Population size = {}".format(len(hps)))
        explore_only = bool(self.do_explore and not self.do_exploit)
        slices = partition(len(hps), len(self.worker_ranks))
        reqs = []
        for rank, sl in zip(self.worker_ranks, slices):
            for cid in sl:
                self.cluster_id_to_worker_rank[cid] = rank
            reqs.append(self.comm.isend(
                protocol.add_graphs([hps[i] for i in sl], sl.start, explore_only), dest=rank))
        for r in reqs:
            r.wait()

    def kill_all_workers(self) -> None:
        self._broadcast(protocol.exit_())

    # ------------------------------------------------------------------ rounds
    @property
    def board_mode(self) -> bool:
        # the file path needs the GET barrier: its checkpoints must be on disk before the master copies them
        return self.scoreboard is not None and self.do_exploit and self.exploit_path != "file"

    def train(self, round_num: int) -> float:
        start = time.time()
        train_sent_ahead = False
        if self.board_mode and self._known is None:
            self.get_all_values()
        for rnd in range(round_num):
            t0 = time.time()
            self._say("\nRound {}".format(rnd))
            if not train_sent_ahead:
                self._broadcast(protocol.train(self.epochs_per_round, self.epochs_per_round * round_num))
            train_sent_ahead = False
            if self.do_exploit:
                with nvtx.range(f"master exploit round {rnd}"):
                    self.exploit()
            if self.do_explore:
                self.explore()
            if self.board_mode:
                # Refresh the master's view of hyper-parameters and membership.  EXPLORE, GET and the next TRAIN
                # go out back to back: the workers answer the GET right after exploring and start the next round
                # without waiting for the master to have read the replies.
                self._broadcast(protocol.get())
                if rnd + 1 < round_num:
                    self._broadcast(protocol.train(self.epochs_per_round, self.epochs_per_round * round_num))
                    train_sent_ahead = True
                self._remember(self._gather())
            self.rounds_done += 1
            self._say("Round elapsed time: {}\n".format(datetime.timedelta(seconds=time.time() - t0)))
        self.flush_all_instructions()
        elapsed = time.time() - start
        self._say("Total elapsed time: {}".format(datetime.timedelta(seconds=elapsed)))
        return elapsed

    def _remember(self, replies: Dict[int, Any]) -> List[List[Any]]:
        """Digest GET replies: who is alive, who owns whom, current hparams."""
        out: List[List[Any]] = []
        owner: Dict[int, int] = {}
        for rank in self.worker_ranks:
            for rec in replies[rank]:
                out.append(list(rec))
                owner[rec[0]] = rank
        self.cluster_id_to_worker_rank = owner
        self._known = {rec[0]: rec for rec in out}
        self.pop_size = len(out)              # the GET view is authoritative: dropped members shrink the population
        return out

    def _gather_values(self) -> List[List[Any]]:
        """The reference's gather: GET to everyone, replies in rank order; doubles as the round barrier."""
        self._broadcast(protocol.get())
        return self._remember(self._gather())

    def _values_from_board(self) -> Optional[List[List[Any]]]:
        """Board path of the ranking input: wait on the device board (the round barrier), take scores from it
        and hyper-parameters from the last GET view.  Returns None when the board timed out."""
        ids = sorted(self._known)
        try:
            scores = self.scoreboard.read(self.rounds_done + 1, members=ids, timeout=self.board_timeout)
        except TimeoutError as e:
            self._say("score board timed out ({}); falling back to the GET gather".format(e))
            self.board_fallbacks += 1
            return None
        self.board_scores = dict(scores)
        out = []
        for cid in ids:
            sc = scores[cid]
            if isinstance(sc, float) and math.isnan(sc):
                continue                          # dropped marker: the worker removes this member
            rec = self._known[cid]
            out.append([cid, sc, copy.deepcopy(rec[2])])
        return out

    def exploit(self) -> None:
        t_gather = time.time()
        all_values: Optional[List[List[Any]]] = None
        if self.board_mode and self._known is not None:
            all_values = self._values_from_board()
        if all_values is None:
            all_values = self._gather_values()
        owner = self.cluster_id_to_worker_rank
        self.gather_time += time.time() - t_gather

        t0 = time.time()
        all_values.sort(key=lambda v: v[1])
        self.pop_size = len(all_values)       # dropped members shrink the population
        pairs = exploit_pairs(all_values)
        per_rank: Dict[int, List[List[Any]]] = {}
        round_log = []
        for top, bottom in pairs:
            winner, loser = all_values[top], all_values[bottom]
            winner_id, loser_id = winner[0], loser[0]
            loser[1] = winner[1]
            loser[2] = copy.deepcopy(winner[2])
            extras: Dict[str, Any] = {"path": self.exploit_path, "src_id": winner_id,
                                      "src_rank": owner[winner_id], "round": self.rounds_done,
                                      "explore_follows": bool(self.do_explore)}
            if self.exploit_path == "file":
                self.exploit_bytes += self.copyfiles(self.member_dir(winner_id), self.member_dir(loser_id))
            per_rank.setdefault(owner[loser_id], []).append([loser_id, loser[1], loser[2], extras])
            if self.exploit_path in TWO_SIDED_PATHS and winner_id != loser_id:
                donor = {"path": self.exploit_path, "role": "src", "dst_id": loser_id,
                         "dst_rank": owner[loser_id], "round": self.rounds_done}
                per_rank.setdefault(owner[winner_id], []).append([winner_id, winner[1], None, donor])
            round_log.append((winner_id, loser_id))
            self._say("Copied: {} -> {}".format(winner_id, loser_id))
        reqs = [self.comm.isend(protocol.set_values(vals), dest=rank)
                for rank, vals in per_rank.items() if rank != self.master_rank]
        for r in reqs:
            r.wait()
        self.history.append({"round": self.rounds_done, "pairs": round_log,
                             "ranking": [v[0] for v in all_values]})
        self.exploit_time += time.time() - t0

    def member_dir(self, cluster_id: int) -> str:
        return os.path.join(self.save_root, "model_" + str(cluster_id))

    def copyfiles(self, src_dir: str, dest_dir: str) -> int:
        if os.path.abspath(src_dir) == os.path.abspath(dest_dir):
            self._say("Warning, src_dir and dest_dir are the same")
            return 0
        return ckpt.copy_member_files(src_dir, dest_dir)

    def explore(self) -> None:
        self._broadcast(protocol.explore())

    def flush_all_instructions(self) -> None:
        # GET blocks until every worker has drained its instruction queue.
        self.get_all_values()

    def get_all_values(self) -> List[List[Any]]:
        return self._gather_values()

    # ------------------------------------------------------------------ reports
    def profiling_info(self) -> Dict[str, float]:
        self._broadcast(protocol.get_profiling_info())
        infos = self._gather()
        vals = [infos[r] for r in self.worker_ranks]
        n = max(len(vals), 1)
        out = {"train": sum(v[0] for v in vals) / n,
               "explore": sum(v[1] for v in vals) / n,
               "exploit": self.exploit_time,
               "gather_wait": self.gather_time,
               "exploit_bytes": float(self.exploit_bytes)}
        # optional third element: per-phase device timers reported by the worker
        extra = [v[2] for v in vals if len(v) > 2 and isinstance(v[2], dict)]
        for key in sorted({k for e in extra for k in e}):
            out["device_" + key] = sum(e.get(key, 0.0) for e in extra) / max(len(extra), 1)
            out["device_" + key + "_max"] = max(e.get(key, 0.0) for e in extra)      # slowest worker
        return out

    def print_profiling_info(self) -> Dict[str, float]:
        info = self.profiling_info()
        self._say("")
        self._say("=======Profiling Information========")
        self._say("Total train time: {}".format(datetime.timedelta(seconds=info["train"])))
        self._say("Total exploit time: {}".format(datetime.timedelta(seconds=info["exploit"])))
        self._say("Total explore time: {}\n".format(datetime.timedelta(seconds=info["explore"])))
        return info

    def dump_all_models_to_json(self, filename: str) -> None:
        report.dump_population(filename, self.get_all_values())
        self._say("Saving all models to {}".format(filename))

    def report_best_model(self) -> Dict[str, Any]:
        path = os.path.join(self.save_root, "best_model.json")
        doc = report.dump_best(path, self.get_all_values())
        self._say("Saving best model to {}".format(path))
        return doc

    def _reports(self, toy: bool = False) -> Dict[str, Optional[str]]:
        out = report.write_reports(self.save_root, self.do_exploit, self.do_explore, toy=toy)
        for k, v in out.items():
            if v:
                self._say("Writing results to {}".format(v))
        return out

    def report_plot_for_toy_model(self):
        return self._reports(toy=True).get("toy")

    def report_accuracy_plot(self):
        return self._reports().get("acc")

    def report_lr_plot(self):
        return self._reports().get("lr")

    def report_best3_plot(self):
        return self._reports().get("best3")
