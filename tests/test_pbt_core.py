"""PBT control logic over the in-process fabric (SURVEY §4.4 items 1 and the NaN-drop rule)."""
import math
import random
import threading
import time

import pytest

from distributedtf_b200 import protocol
from distributedtf_b200.master import PBTCluster, exploit_pairs, partition
from distributedtf_b200.model_base import ModelBase
from distributedtf_b200.parallel.local import LocalFabric
from distributedtf_b200.protocol import WorkerInstruction as WI
from distributedtf_b200.worker import TrainingWorker


class ScoreModel(ModelBase):
    """Deterministic member: accuracy = base(id) + epochs; records every call."""
    log = []
    lock = threading.Lock()

    def __init__(self, cluster_id, hparams, save_base_dir, **kw):
        super().__init__(cluster_id, hparams, save_base_dir)
        self.adopted = []

    def train(self, num_epoch, total_epochs):
        self.epoches_trained += num_epoch
        self.accuracy = float(self.cluster_id) + 0.001 * self.epoches_trained
        with ScoreModel.lock:
            ScoreModel.log.append(("train", self.cluster_id, num_epoch, total_epochs))

    def adopt_state(self, extras):
        self.adopted.append(extras)


def _hp(i):
    return {"opt_case": {"optimizer": "gd", "lr": 0.1}, "batch_size": 100, "decay_rate": round(0.33 + 0.05 * i, 2)}   # never equal to a 1-digit perturbed value


def start_cluster(pop, workers, tmp_path, model=ScoreModel, **kw):
    fabric = LocalFabric(workers + 1)
    ws = [TrainingWorker(fabric.endpoint(r), 0, model, save_base_dir=str(tmp_path / "model_"),
                         verbose=False, fault_hook=kw.pop("fault_hook", None))
          for r in range(1, workers + 1)]
    threads = [threading.Thread(target=w.main_loop, daemon=True) for w in ws]
    for t in threads:
        t.start()
    cluster = PBTCluster(pop, fabric.endpoint(0), 0, kw.pop("epochs_per_round", 1),
                         save_root=str(tmp_path), verbose=False,
                         initial_hparams=[_hp(i) for i in range(pop)], **kw)
    return cluster, ws, threads, fabric


@pytest.mark.parametrize("pop,workers", [(1, 1), (2, 1), (5, 2), (7, 3), (8, 8), (3, 5), (20, 4), (40, 4), (9, 4)])
def test_partition_matches_reference_rule(pop, workers):
    parts = partition(pop, workers)
    per = math.ceil(pop / workers)
    flat = [i for p in parts for i in p]
    assert flat == list(range(pop))                       # contiguous, complete, disjoint
    assert all(len(p) <= per for p in parts)
    assert [p.start for p in parts if len(p)] == [k * per for k in range(len([p for p in parts if len(p)]))]


@pytest.mark.parametrize("pop", list(range(1, 10)) + [20, 40])
def test_exploit_pairing(pop):
    pairs = exploit_pairs(list(range(pop)))
    n = math.ceil(pop / 4)
    assert len(pairs) == n
    assert [b for _, b in pairs] == list(range(n))                 # bottom n, ascending
    assert [t for t, _ in pairs] == list(range(pop - n, pop))      # top n, same order


@pytest.mark.parametrize("pop,workers", [(4, 1), (8, 3), (9, 4)])
def test_round_exploit_explore(pop, workers, tmp_path):
    ScoreModel.log.clear()
    cluster, ws, threads, fabric = start_cluster(pop, workers, tmp_path, exploit_path="p2p")
    cluster.train(2)
    values = sorted(cluster.get_all_values())
    cluster.kill_all_workers()
    for t in threads:
        t.join(5)
    assert len(values) == pop
    n = math.ceil(pop / 4)
    # ranking is by id (accuracy = id + eps): losers 0..n-1 got the hparams of winners pop-n..pop-1
    for h in cluster.history:
        assert h["pairs"] == [(pop - n + i, i) for i in range(n)]
    members = {g.cluster_id: g for w in ws for g in w.worker_graphs}
    for i in range(n):
        loser = members[i]
        assert [e["src_id"] for e in loser.adopted] == [pop - n + i] * 2
        assert loser.adopted[0]["src_rank"] == cluster.cluster_id_to_worker_rank[pop - n + i]
        assert loser.need_explore is False          # explored right after SET
        assert loser.hparams["decay_rate"] != _hp(i)["decay_rate"]     # adopted + perturbed winner value
    untouched = members[pop - 1]
    assert untouched.hparams == _hp(pop - 1) and untouched.adopted == []
    # every member trained once per round with (epochs, epochs*rounds)
    assert sorted(ScoreModel.log) == sorted([("train", i, 1, 2) for i in range(pop)] * 2)


@pytest.mark.parametrize("do_exploit,do_explore", [(True, True), (True, False), (False, True), (False, False)])
def test_mode_matrix(do_exploit, do_explore, tmp_path):
    pop = 4
    cluster, ws, threads, _ = start_cluster(pop, 2, tmp_path, do_exploit=do_exploit,
                                            do_explore=do_explore, exploit_path="p2p")
    cluster.train(1)
    cluster.kill_all_workers()
    for t in threads:
        t.join(5)
    members = {g.cluster_id: g for w in ws for g in w.worker_graphs}
    changed = {i for i, m in members.items() if m.hparams != _hp(i)}
    assert all(w.is_expolore_only == (do_explore and not do_exploit) for w in ws)
    if do_exploit and do_explore:
        assert changed == {0}
    elif do_exploit:
        assert changed == {0} and members[0].hparams == _hp(3)   # copied, not perturbed
        assert members[0].need_explore is True
    elif do_explore:
        assert changed == {0, 1, 2, 3}                           # explore-only perturbs everyone
    else:
        assert changed == set()


def test_get_is_a_barrier_and_instructions_are_fifo(tmp_path):
    order = []

    class Slow(ScoreModel):
        def train(self, n, t):
            time.sleep(0.05)
            super().train(n, t)
            order.append(("train", self.cluster_id))

        def perturb_hparams(self):
            order.append(("explore", self.cluster_id))

    cluster, ws, threads, _ = start_cluster(2, 1, tmp_path, model=Slow, do_exploit=False)
    cluster.train(2)
    order.append("flushed")
    cluster.kill_all_workers()
    for t in threads:
        t.join(5)
    assert order == [("train", 0), ("train", 1), ("explore", 0), ("explore", 1)] * 2 + ["flushed"]


def test_member_drop_on_exception_and_nan(tmp_path):
    class Flaky(ScoreModel):
        def train(self, n, t):
            super().train(n, t)
            if self.cluster_id == 1 and self.epoches_trained == 2:
                raise RuntimeError("boom")
            if self.cluster_id == 2:
                self.accuracy = float("nan")

    cluster, ws, threads, _ = start_cluster(6, 2, tmp_path, model=Flaky, exploit_path="p2p")
    for i in range(6):
        (tmp_path / f"model_{i}").mkdir()
    cluster.train(2)
    ids = sorted(v[0] for v in cluster.get_all_values())
    cluster.kill_all_workers()
    for t in threads:
        t.join(5)
    assert ids == [0, 3, 4, 5]
    assert cluster.pop_size == 4                         # refreshed from the gathered list
    assert not (tmp_path / "model_1").exists() and not (tmp_path / "model_2").exists()
    assert sorted(d for w in ws for d in w.dropped) == [1, 2]
    # pairs never name a dropped member
    assert all(a in ids or a in (1,) for h in cluster.history[1:] for a, b in h["pairs"])
    assert all(x not in (1, 2) for x in cluster.history[-1]["ranking"])


def test_injected_fault_hook(tmp_path):
    def hook(member, rnd):
        if member.cluster_id == 0 and rnd == 1:
            raise FloatingPointError("injected")
    cluster, ws, threads, _ = start_cluster(3, 1, tmp_path, fault_hook=hook, exploit_path="p2p")
    cluster.train(3)
    ids = sorted(v[0] for v in cluster.get_all_values())
    cluster.kill_all_workers()
    assert ids == [1, 2]


def test_profiling_info_is_mean_over_workers(tmp_path):
    cluster, ws, threads, _ = start_cluster(4, 2, tmp_path, exploit_path="p2p")
    cluster.train(1)
    info = cluster.print_profiling_info()
    cluster.kill_all_workers()
    for t in threads:
        t.join(5)
    assert info["train"] == pytest.approx(sum(w.train_time for w in ws) / 2)
    assert info["explore"] == pytest.approx(sum(w.explore_time for w in ws) / 2)
    assert info["exploit"] >= 0 and info["gather_wait"] >= 0


def test_two_sided_paths_send_donor_records(tmp_path):
    donated = []

    class Donor(ScoreModel):
        def donate_state(self, extras):
            donated.append((self.cluster_id, extras["dst_id"], extras["dst_rank"]))

    cluster, ws, threads, _ = start_cluster(8, 4, tmp_path, model=Donor, exploit_path="nccl")
    cluster.train(1)
    cluster.kill_all_workers()
    for t in threads:
        t.join(5)
    assert sorted(donated) == [(6, 0, 1), (7, 1, 1)]
    members = {g.cluster_id: g for w in ws for g in w.worker_graphs}
    assert members[6].hparams == _hp(6) and members[6].need_explore is False


def test_master_never_sends_to_itself(tmp_path):
    cluster, ws, threads, fabric = start_cluster(4, 2, tmp_path, exploit_path="p2p")
    cluster.train(1)
    cluster.kill_all_workers()
    assert fabric._queues[(0, 0)].empty()


def test_recv_timeout_detects_dead_worker(tmp_path):
    from distributedtf_b200.parallel.comm_base import CommTimeout
    fabric = LocalFabric(2)     # no worker thread started: the worker is "dead"
    with pytest.raises(CommTimeout):
        c = PBTCluster(2, fabric.endpoint(0), 0, 1, save_root=str(tmp_path), verbose=False,
                       recv_timeout=0.2, initial_hparams=[_hp(0), _hp(1)])
        c.get_all_values()


def test_members_in_flight_matches_sequential_on_toy(tmp_path):
    """Thread-pool schedule of a worker's members (host members: no streams involved) gives the same run."""
    from distributedtf_b200.config import RunConfig
    from distributedtf_b200.launch import run_local
    outs = []
    for k in (1, 3):
        cfg = RunConfig(target_model="toy", population_size=6, train_round=4, epochs_per_round=2, workers=2,
                        transport="local", save_root=str(tmp_path / f"sd{k}"), results_file=str(tmp_path / f"r{k}.txt"),
                        verbose=False, seed=11, members_in_flight=k).validate()
        outs.append(run_local(cfg))
    assert outs[0]["best"] == outs[1]["best"]
    assert [h["pairs"] for h in outs[0]["history"]] == [h["pairs"] for h in outs[1]["history"]]


# --------------------------------------------------------------------------- device score board as the barrier
class FakeBoard:
    """Stands in for parallel/scoreboard.ScoreBoard: members write (score, epoch) into a dict."""

    def __init__(self):
        self.slots = {}
        self.lock = threading.Lock()
        self.reads = []

    def put(self, member, score, epoch):
        with self.lock:
            self.slots[member] = (score, epoch)

    def read(self, epoch, members=None, timeout=60.0):
        deadline = time.time() + timeout
        ids = list(members)
        while True:
            with self.lock:
                if all(self.slots.get(i, (0, 0))[1] >= epoch for i in ids):
                    self.reads.append((epoch, tuple(ids)))
                    return {i: self.slots[i][0] for i in ids}
            if time.time() > deadline:
                raise TimeoutError("members did not publish")
            time.sleep(0.001)


class BoardModel(ScoreModel):
    """Publishes its score on the fake board at the end of train, like a GPU member's evaluation kernels."""
    board = None
    silent = set()          # members that never publish (board time-out path)
    nan_at = {}             # member -> round at which its score turns NaN
    seen = []               # instruction trace of the workers

    def train(self, num_epoch, total_epochs):
        super().train(num_epoch, total_epochs)
        self.rounds = getattr(self, "rounds", 0) + 1
        if BoardModel.nan_at.get(self.cluster_id) == self.rounds:
            self.accuracy = float("nan")
        if self.cluster_id not in BoardModel.silent:
            BoardModel.board.put(self.cluster_id, self.accuracy, self.rounds)


class TracingWorker(TrainingWorker):
    def main_loop(self):
        orig = self.comm.recv

        def recv(source, **kw):
            msg = orig(source=source, **kw)
            BoardModel.seen.append((self.rank, WI(msg[0]).name))
            return msg
        self.comm.recv = recv
        super().main_loop()


def _board_cluster(pop, workers, tmp_path, **kw):
    BoardModel.board, BoardModel.silent, BoardModel.nan_at, BoardModel.seen = FakeBoard(), set(), {}, []
    fabric = LocalFabric(workers + 1)
    ws = [TracingWorker(fabric.endpoint(r), 0, BoardModel, save_base_dir=str(tmp_path / "model_"), verbose=False)
          for r in range(1, workers + 1)]
    threads = [threading.Thread(target=w.main_loop, daemon=True) for w in ws]
    for t in threads:
        t.start()
    cluster = PBTCluster(pop, fabric.endpoint(0), 0, 1, save_root=str(tmp_path), verbose=False, exploit_path="p2p",
                         initial_hparams=[_hp(i) for i in range(pop)], scoreboard=BoardModel.board, **kw)
    return cluster, ws, threads


def test_board_is_the_round_barrier_and_get_leaves_the_critical_path(tmp_path):
    """With a score board the master ranks without a GET between TRAIN and SET; the hyper-parameter refresh
    (GET) and the next TRAIN follow EXPLORE back to back."""
    pop = 8
    cluster, ws, threads = _board_cluster(pop, 2, tmp_path)
    cluster.train(3)
    values = sorted(cluster.get_all_values())
    cluster.kill_all_workers()
    for t in threads:
        t.join(5)
    assert cluster.board_fallbacks == 0 and [e for e, _ in BoardModel.board.reads] == [1, 2, 3]
    trace = [name for rank, name in BoardModel.seen if rank == 1]
    # initial view, then per round TRAIN -> SET (rank 1 owns the losers) -> EXPLORE -> GET, flush GET, our GET, EXIT
    assert trace == ["ADD_GRAPHS", "GET",
                     "TRAIN", "SET", "EXPLORE", "GET",
                     "TRAIN", "SET", "EXPLORE", "GET",
                     "TRAIN", "SET", "EXPLORE", "GET", "GET", "GET", "EXIT"]
    n = math.ceil(pop / 4)
    for h in cluster.history:           # same pairing as the GET path (scores = id + eps)
        assert h["pairs"] == [(pop - n + i, i) for i in range(n)]
    members = {g.cluster_id: g for w in ws for g in w.worker_graphs}
    assert [e["src_id"] for e in members[0].adopted] == [pop - n] * 3
    # the master's view follows the explored hyper-parameters of the losers
    assert {v[0]: v[2] for v in values} == {i: m.hparams for i, m in members.items()}


def test_board_dropped_marker_and_timeout_fallback(tmp_path):
    cluster, ws, threads = _board_cluster(6, 2, tmp_path)
    BoardModel.nan_at = {5: 1}            # best member turns NaN in round 1: marker on the board, worker drops it
    cluster.train(1)
    assert cluster.pop_size == 5 and cluster.history[0]["pairs"] == [(3, 0), (4, 1)]
    BoardModel.silent = {2}               # a member that never reaches the board: time out, rank from a GET gather
    cluster.board_timeout = 0.3
    cluster.train(1)
    assert cluster.board_fallbacks == 1 and len(cluster.history) == 2
    values = cluster.get_all_values()
    cluster.kill_all_workers()
    for t in threads:
        t.join(5)
    assert sorted(v[0] for v in values) == [0, 1, 2, 3, 4]
