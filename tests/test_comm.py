"""Control-plane transports: ordering, isolation by source, multi-process TCP and gloo."""
import multiprocessing as mp
import os
import socket
import threading

import pytest

from distributedtf_b200.parallel.comm_base import CommTimeout
from distributedtf_b200.parallel.local import LocalFabric


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_local_fifo_per_pair_and_copy_semantics():
    f = LocalFabric(3)
    m, w1, w2 = f.endpoint(0), f.endpoint(1), f.endpoint(2)
    payload = {"k": [1, 2]}
    for i in range(5):
        m.isend((i, payload), dest=1).wait()
    w2.send("from2", dest=0)
    w1.send("from1", dest=0)
    got = [w1.recv(source=0) for _ in range(5)]
    assert [g[0] for g in got] == list(range(5))
    got[0][1]["k"].append(3)
    assert payload == {"k": [1, 2]}
    assert m.recv(source=1) == "from1" and m.recv(source=2) == "from2"
    with pytest.raises(CommTimeout):
        m.recv(source=1, timeout=0.05)
    assert (m.Get_rank(), m.Get_size(), w2.rank, w2.size) == (0, 3, 2, 3)


def _tcp_worker(port, rank, size):
    from distributedtf_b200.parallel.tcp import TcpWorkerComm
    c = TcpWorkerComm("127.0.0.1", port, rank, size)
    while True:
        msg = c.recv(source=0)
        if msg == "stop":
            break
        c.send((rank, msg), dest=0)
    c.close()


def test_tcp_star_multi_process():
    from distributedtf_b200.parallel.tcp import TcpMasterComm
    size = 4
    master = TcpMasterComm("127.0.0.1", 0, size)
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_tcp_worker, args=(master.port, r, size)) for r in range(1, size)]
    for p in procs:
        p.start()
    master.wait_for_workers()
    for i in range(20):
        reqs = [master.isend({"i": i, "blob": b"x" * (1000 * i)}, dest=r) for r in range(1, size)]
        for r in reqs:
            r.wait()
    for r in range(size - 1, 0, -1):        # receive in reverse rank order on purpose
        for i in range(20):
            src, msg = master.recv(source=r, timeout=30)
            assert src == r and msg["i"] == i and len(msg["blob"]) == 1000 * i
    for r in range(1, size):
        master.send("stop", dest=r)
    for p in procs:
        p.join(20)
        assert p.exitcode == 0
    master.close()


def _gloo_rank(rank, size, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(size))
    import torch.distributed as dist
    from distributedtf_b200.parallel.gloo import GlooComm
    dist.init_process_group("gloo", rank=rank, world_size=size)
    c = GlooComm()
    if rank == 0:
        for r in range(1, size):
            c.isend(("hello", r, list(range(100))), dest=r).wait()
        q.put([c.recv(source=r) for r in range(1, size)])
    else:
        msg = c.recv(source=0)
        c.send((rank, msg[1], sum(msg[2])), dest=0)
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_comm_world_size_3():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_rank, args=(r, 3, port, q)) for r in range(3)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert out == [(1, 1, 4950), (2, 2, 4950)]


def test_toy_pbt_over_tcp_processes(tmp_path):
    """BASELINE config 1: toy PBT, 1 master + 1 worker as separate processes."""
    from distributedtf_b200.config import RunConfig
    from distributedtf_b200.launch import spawn_tcp
    cfg = RunConfig(target_model="toy", population_size=2, train_round=3, epochs_per_round=4,
                    workers=1, transport="tcp", master_port=0, save_root=str(tmp_path / "sd"),
                    results_file=str(tmp_path / "res.txt"), verbose=False, seed=1).validate()
    out = spawn_tcp(cfg)
    assert out["pop_size"] == 2
    assert (tmp_path / "sd" / "best_model.json").exists()
    assert "n = 2, pop_size = 2, time = " in (tmp_path / "res.txt").read_text()


def test_local_fabric_abort_unblocks_recv():
    """A dead endpoint must not leave its peers blocked forever (run_local aborts the fabric on a worker error)."""
    import threading
    from distributedtf_b200.parallel.local import LocalFabric
    fabric = LocalFabric(2)
    got = []

    def waiter():
        try:
            fabric.endpoint(0).recv(source=1)
        except RuntimeError as e:
            got.append(e)

    t = threading.Thread(target=waiter, daemon=True)
    t.start()
    fabric.abort(ValueError("worker died"))
    t.join(5)
    assert not t.is_alive() and got and "worker died" in str(got[0])


def test_tcp_master_notices_a_dead_worker():
    """a worker process that dies must surface at the master as an error, not as a hang."""
    import socket
    import threading
    from distributedtf_b200.parallel.tcp import TcpMasterComm, TcpWorkerComm
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    master = TcpMasterComm("127.0.0.1", port, 2)
    worker = TcpWorkerComm("127.0.0.1", port, 1, 2)
    master.wait_for_workers()
    master.isend(("ping",), dest=1)
    assert worker.recv(source=0, timeout=10) == ("ping",)
    worker.close()                                   # the worker goes away without answering
    with pytest.raises(ConnectionError):
        master.recv(source=1, timeout=10)
    master.close()


def _gloo_silent_rank(rank, size, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(size))
    import torch.distributed as dist
    from distributedtf_b200.parallel.comm_base import CommTimeout
    from distributedtf_b200.parallel.gloo import GlooComm
    dist.init_process_group("gloo", rank=rank, world_size=size)
    c = GlooComm()
    if rank == 0:
        try:
            c.recv(source=1, timeout=1.0)        # the worker never answers
            q.put("no timeout")
        except CommTimeout as e:
            q.put("timeout: " + str(e))
        c.isend("done", dest=1).wait()
    else:
        assert c.recv(source=0, timeout=60) == "done"
    dist.destroy_process_group()


def test_gloo_recv_honours_timeout():
    """A silent worker surfaces as CommTimeout on the gloo transport too (it used to block forever)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_silent_rank, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(60)
    assert out.startswith("timeout:"), out


def test_mpi_adapter_against_a_fake_mpi4py(monkeypatch):
    """mpi4py is not on the image: drive MpiComm through a stand-in COMM_WORLD (queues between two threads)
    so that its isend / recv / timeout code has run at least once."""
    import queue
    import sys
    import threading
    import types
    from distributedtf_b200.parallel.comm_base import CommTimeout
    boxes = {(s, d): queue.Queue() for s in range(2) for d in range(2)}
    local = threading.local()

    class _Req:
        def wait(self):
            return None

    class _World:
        def Get_rank(self):
            return local.rank

        def Get_size(self):
            return 2

        def isend(self, obj, dest):
            boxes[(local.rank, dest)].put(obj)
            return _Req()

        def iprobe(self, source):
            return not boxes[(source, local.rank)].empty()

        def recv(self, source):
            return boxes[(source, local.rank)].get()

    fake = types.ModuleType("mpi4py")
    fake.MPI = types.SimpleNamespace(COMM_WORLD=_World())
    monkeypatch.setitem(sys.modules, "mpi4py", fake)
    from distributedtf_b200.parallel import mpi
    assert mpi.available()
    out = {}

    def worker():
        local.rank = 1
        c = mpi.MpiComm()
        msg = c.recv(source=0, timeout=10)
        c.send(("echo", msg, c.Get_rank(), c.Get_size()), dest=0)

    t = threading.Thread(target=worker, daemon=True)
    t.start()
    local.rank = 0
    c = mpi.MpiComm()
    c.isend({"k": 1}, dest=1).wait()
    out = c.recv(source=1, timeout=10)
    t.join(5)
    assert out == ("echo", {"k": 1}, 1, 2)
    with pytest.raises(CommTimeout):
        c.recv(source=1, timeout=0.05)
