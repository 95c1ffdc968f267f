"""Entry point and role split.

Parity: reference ``main_manager.py:46-73`` — every endpoint runs the same
program; the master wipes/creates ``savedata``, builds the ``PBTCluster``, dumps
the initial hyper-parameters, trains ``train_round`` rounds, appends the timing
line to ``test_results.txt``, writes reports and the best model, prints the
profiling summary and sends EXIT; every other endpoint runs
``TrainingWorker.main_loop()``.

Launch modes (no ``mpirun`` on the target image):

* ``transport="local"``  master + workers as threads of this process (tests, 1 GPU);
* ``transport="tcp"``    ``spawn_tcp`` starts one process per worker (one per GPU)
  and runs the master in the parent; also used under ``torchrun`` where local
  rank 0 hosts the master thread next to its worker;
* ``transport="gloo"`` / ``"mpi"``  the process is already one rank of a job.
"""
from __future__ import annotations

import json
import multiprocessing as mp
import os
import random
import shutil
import threading
from typing import Any, Callable, Dict, List, Optional

from . import report
from .config import RunConfig, parse_args
from .master import PBTCluster
from .models import get_model_class
from .parallel.local import LocalFabric
from .worker import TrainingWorker


def _fault_hook(cfg: RunConfig) -> Optional[Callable[[Any, int], None]]:
    plan = {}
    for item in cfg.inject_nan:
        mid, rnd = item.split("@")
        plan.setdefault(int(rnd), set()).add(int(mid))
    if not plan:
        return None

    def hook(member, rnd):
        if member.cluster_id in plan.get(rnd, ()):
            raise FloatingPointError(f"injected NaN for member {member.cluster_id} at round {rnd}")
    return hook


def model_kwargs(cfg: RunConfig, dp_group=None, dp_rank: int = 0) -> Dict[str, Any]:
    if cfg.target_model == "toy":
        return {}
    kw = dict(engine=cfg.engine, dtype=cfg.dtype, synthetic=cfg.synthetic, data_dir=cfg.data_dir,
              max_train_steps=cfg.max_train_steps, eval_batches=cfg.eval_batches,
              resume=cfg.resume)
    kw.update(epochs_between_evals=cfg.epochs_between_evals, stop_threshold=cfg.stop_threshold,
              export_dir=cfg.export_dir)
    if cfg.target_model == "cifar10":
        kw.update(resnet_size=cfg.resnet_size, resnet_version=cfg.resnet_version, loss_scale=cfg.loss_scale,
                  resnet_bottleneck=cfg.resnet_bottleneck)
    # reference distribution_utils.get_distribution_strategy: one device, or a mirrored (data-parallel) member
    from .utils.misc.distribution_utils import get_distribution_strategy
    strategy = get_distribution_strategy(cfg.dp_size if cfg.target_model != "toy" else 1, cfg.all_reduce_alg)
    if strategy.kind == "data_parallel":
        kw.update(dp_group=dp_group, dp_rank=dp_rank, dp_size=strategy.num_gpus,
                  all_reduce_alg=strategy.all_reduce_alg)
    if getattr(cfg, "_board_info", None):
        kw["scoreboard"] = cfg._board_info
    if cfg.hooks and cfg.verbose:
        from .utils.logs import hooks as _hooks
        kw["hooks"] = _hooks.get_train_hooks(cfg.hooks, model_dir=cfg.save_root, batch_size=128)
    # Members that train concurrently inside one process (members_in_flight > 1: a host thread and a CUDA stream
    # each) replay CUDA graphs like everybody else.  For a while they could not: captures were invalidated or ended
    # in illegal memory accesses in four of five runs.  The cause was torch.cuda.Stream() handing out the same 32
    # pooled streams round-robin, so one member's capture stream was another member's compute stream; every
    # long-lived stream now comes from utils/streams.new_stream().  DTF_INFLIGHT_GRAPHS=0 restores direct launches.
    kw["use_graphs"] = cfg.members_in_flight <= 1 or os.environ.get("DTF_INFLIGHT_GRAPHS", "1") != "0"
    kw["benchmark"] = (cfg.benchmark_logger_type, cfg.benchmark_log_dir)
    # the file-exploit path copies checkpoint files right after GET: keep those saves synchronous
    kw["checkpoint_every"] = cfg.checkpoint_every
    kw["async_checkpoint"] = cfg.async_checkpoint if cfg.async_checkpoint is not None else cfg.exploit_path != "file"
    return kw


def make_scoreboard(cfg: RunConfig, same_process: bool):
    """Device score board owned by the master.  Returns (board, info-for-members)."""
    if not cfg.scoreboard or cfg.target_model == "toy":
        return None, None
    import torch
    if not torch.cuda.is_available():
        return None, None
    from .parallel.scoreboard import ScoreBoard
    board = ScoreBoard(max(cfg.population_size, 1), torch.device("cuda", torch.cuda.current_device()))
    # members in the owner's process use the raw pointer, everyone else maps the CUDA IPC handle
    return board, {"ptr": board.ptr, "owner_pid": os.getpid(), "handle": board.handle()}


def run_master(cfg: RunConfig, comm, board=None) -> Dict[str, Any]:
    # main_manager.py:49-50 wipes savedata unconditionally; here `clean` can be switched off (resume)
    from types import SimpleNamespace
    from .utils.misc.model_helpers import apply_clean
    apply_clean(SimpleNamespace(clean=cfg.clean, model_dir=cfg.save_root))
    os.makedirs(cfg.save_root, exist_ok=True)
    initial = None
    if cfg.initial_hp:
        with open(cfg.initial_hp) as f:
            initial = [rec["hparams"] for rec in json.load(f)]
    rng = random.Random(cfg.seed) if cfg.seed is not None else None
    cluster = PBTCluster(cfg.population_size, comm, cfg.master_rank,
                         epochs_per_round=cfg.epochs_per_round,
                         do_exploit=cfg.do_exploit, do_explore=cfg.do_explore,
                         save_root=cfg.save_root, exploit_path=cfg.exploit_path, rng=rng,
                         recv_timeout=cfg.recv_timeout, initial_hparams=initial,
                         scoreboard=board, verbose=cfg.verbose)
    cluster.dump_all_models_to_json(os.path.join(cfg.save_root, "initial_hp.json"))
    elapsed = cluster.train(cfg.train_round)
    if cfg.results_file:
        report.append_timing(cfg.results_file, comm.Get_size(), cfg.population_size, elapsed)
    if cfg.target_model == "toy":
        cluster.report_plot_for_toy_model()
    else:
        cluster._reports()
    best = cluster.report_best_model()
    prof = cluster.print_profiling_info()
    cluster.kill_all_workers()
    return {"elapsed": elapsed, "best": best, "profiling": prof, "history": cluster.history,
            "pop_size": cluster.pop_size, "board_fallbacks": cluster.board_fallbacks}


def run_worker(cfg: RunConfig, comm, dp_group=None, mirror=None) -> TrainingWorker:
    worker = TrainingWorker(comm, cfg.master_rank, get_model_class(cfg.target_model),
                            save_base_dir=os.path.join(cfg.save_root, "model_"),
                            verbose=cfg.verbose, model_kwargs=model_kwargs(cfg, dp_group, 0),
                            fault_hook=_fault_hook(cfg), mirror=mirror,
                            members_in_flight=cfg.members_in_flight, sm_limit=cfg.sm_limit)
    if cfg.seed is not None:
        random.seed(cfg.seed * 1000003 + comm.Get_rank())
    worker.main_loop()
    return worker


def run_local(cfg: RunConfig) -> Dict[str, Any]:
    """Master + workers as threads over the in-process fabric."""
    fabric = LocalFabric(cfg.workers + 1)
    board, info = make_scoreboard(cfg, same_process=True)
    object.__setattr__(cfg, "_board_info", info)
    errors: List[BaseException] = []

    def guard(fn, *a):
        try:
            fn(*a)
        except BaseException as e:  # surfaced after join; unblocks the master's recv
            errors.append(e)
            fabric.abort(e)

    threads = [threading.Thread(target=guard, args=(run_worker, cfg, fabric.endpoint(r)),
                                name=f"dtf-worker{r}", daemon=True)
               for r in range(1, cfg.workers + 1)]
    for t in threads:
        t.start()
    try:
        result = run_master(cfg, fabric.endpoint(0), board)
    except RuntimeError:
        if errors:
            raise errors[0]
        raise
    for t in threads:
        t.join(timeout=60)
    if errors:
        raise errors[0]
    return result


def _init_worker_group(cfg: RunConfig, rank: int, group_port: int, dp_rank: int = 0):
    """Process group among the WORKERS (the master is not in it).  It carries the 64-byte CUDA IPC
    handles of the members' symmetric buffers (device exploit paths) and, for the library baselines,
    the NCCL / gloo state transfers.  Never used for control messages."""
    import torch
    import torch.distributed as dist
    backend = "cpu:gloo,cuda:nccl" if torch.cuda.is_available() else "gloo"
    kw = {}
    if torch.cuda.is_available():
        idx = int(os.environ.get("DTF_LOCAL_GPU", "0")) % torch.cuda.device_count()
        torch.cuda.set_device(idx)
        kw["device_id"] = torch.device("cuda", idx)
    world_rank = (rank - 1) * cfg.dp_size + dp_rank
    os.environ["RANK"] = str(world_rank)
    dist.init_process_group(backend, init_method=f"tcp://{cfg.master_addr}:{group_port}", rank=world_rank,
                            world_size=cfg.workers * cfg.dp_size, **kw)
    dp_group = None
    if cfg.dp_size > 1:            # one sub-group per member group; every process creates all of them
        for w in range(cfg.workers):
            grp = dist.new_group(list(range(w * cfg.dp_size, (w + 1) * cfg.dp_size)))
            if w == rank - 1:
                dp_group = grp
    return dp_group


def _tcp_worker_entry(cfg: RunConfig, rank: int, gpu: Optional[int], group_port: int = 0, dp_rank: int = 0) -> None:
    from .parallel.tcp import TcpWorkerComm
    if gpu is not None:
        os.environ["DTF_LOCAL_GPU"] = str(gpu)
    dp_group = None
    if group_port and (cfg.workers * cfg.dp_size) > 1 and (cfg.exploit_path != "file" or cfg.dp_size > 1):
        dp_group = _init_worker_group(cfg, rank, group_port, dp_rank)
    if dp_rank > 0:                # follower replica: driven by its group leader, never by the master
        import torch.distributed as dist
        from .worker import FollowerWorker
        leader = (rank - 1) * cfg.dp_size

        def recv_from_leader():
            box = [None]
            dist.broadcast_object_list(box, src=leader, group=dp_group)
            return box[0]
        FollowerWorker(recv_from_leader, get_model_class(cfg.target_model),
                       save_base_dir=os.path.join(cfg.save_root, f".replica{dp_rank}_model_"), verbose=False,
                       model_kwargs=model_kwargs(cfg, dp_group, dp_rank)).main_loop()
        dist.destroy_process_group()
        return
    mirror = None
    if cfg.dp_size > 1:
        import torch.distributed as dist
        me = (rank - 1) * cfg.dp_size

        def mirror(msg):
            dist.broadcast_object_list([msg], src=me, group=dp_group)
    comm = TcpWorkerComm(cfg.master_addr, cfg.master_port, rank, cfg.workers + 1)
    try:
        run_worker(cfg, comm, dp_group=dp_group, mirror=mirror)
    except BaseException:
        import traceback
        print(f"[worker {rank}] fatal:", flush=True)
        traceback.print_exc()
        raise
    finally:
        comm.close()
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:
            pass


def spawn_tcp(cfg: RunConfig, gpus: Optional[List[int]] = None) -> Dict[str, Any]:
    """One OS process per worker (one per GPU when ``gpus`` is given), master here."""
    from .parallel.tcp import TcpMasterComm
    ctx = mp.get_context("spawn")
    master = TcpMasterComm(cfg.master_addr, cfg.master_port, cfg.workers + 1)
    cfg = cfg.replace(master_port=master.port)
    board, info = make_scoreboard(cfg, same_process=False)
    object.__setattr__(cfg, "_board_info", info)
    import socket
    with socket.socket() as probe:                       # a free port for the workers' own process group
        probe.bind((cfg.master_addr, 0))
        group_port = probe.getsockname()[1]
    if gpus is None:
        try:
            import torch
            if torch.cuda.is_available() and cfg.target_model != "toy":
                gpus = list(range(torch.cuda.device_count()))
        except ImportError:
            pass
    procs = []
    for r in range(1, cfg.workers + 1):
        for d in range(cfg.dp_size):
            slot = (r - 1) * cfg.dp_size + d
            gpu = gpus[slot % len(gpus)] if gpus else None
            p = ctx.Process(target=_tcp_worker_entry, args=(cfg, r, gpu, group_port, d), daemon=True)
            p.start()
            procs.append(p)
    try:
        master.wait_for_workers()
        result = run_master(cfg, master, board)
    finally:
        for p in procs:
            p.join(timeout=60)
            if p.is_alive():
                p.terminate()
        master.close()
    return result


def run_torchrun(cfg: RunConfig) -> Optional[Dict[str, Any]]:
    """One process per GPU under ``python -m torch.distributed.run``: every rank is a worker endpoint
    (rank + 1); rank 0 additionally hosts the master in a thread.  Control messages go over the TCP
    star on MASTER_PORT + 1; the NCCL/gloo process group carries the symmetric-memory handles."""
    import torch
    import torch.distributed as dist
    from .parallel.tcp import TcpMasterComm, TcpWorkerComm
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    port = int(os.environ.get("MASTER_PORT", "29500")) + 1
    os.environ["DTF_LOCAL_GPU"] = str(local)
    if torch.cuda.is_available():
        torch.cuda.set_device(local % torch.cuda.device_count())
    if not dist.is_initialized():
        dist.init_process_group("cpu:gloo,cuda:nccl" if torch.cuda.is_available() else "gloo")
    cfg = cfg.replace(workers=world, master_addr=addr, master_port=port)
    result: Dict[str, Any] = {}
    master_thread = None
    board, info = (make_scoreboard(cfg, same_process=False) if rank == 0 else (None, None))
    box = [info]
    dist.broadcast_object_list(box, src=0)          # every rank's members need the board's handle
    object.__setattr__(cfg, "_board_info", box[0])
    if rank == 0:
        master = TcpMasterComm(addr, port, world + 1)

        def _m():
            master.wait_for_workers()
            result.update(run_master(cfg, master, board))
        master_thread = threading.Thread(target=_m, name="dtf-master", daemon=True)
        master_thread.start()
    comm = TcpWorkerComm(addr, port, rank + 1, world + 1)
    try:
        run_worker(cfg, comm)
    finally:
        comm.close()
    if master_thread is not None:
        master_thread.join(timeout=120)
    dist.barrier()
    return result if rank == 0 else None


def run(cfg: RunConfig) -> Optional[Dict[str, Any]]:
    cfg.validate()
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ and cfg.transport in ("tcp", "torchrun"):
        return run_torchrun(cfg)
    if cfg.transport == "local":
        return run_local(cfg)
    if cfg.transport == "tcp":
        return spawn_tcp(cfg)
    if cfg.transport == "gloo":
        import torch.distributed as dist
        from .parallel.gloo import GlooComm
        if not dist.is_initialized():
            dist.init_process_group("gloo")
        comm = GlooComm()
    elif cfg.transport == "mpi":
        from .parallel.mpi import MpiComm
        comm = MpiComm()
    else:
        raise ValueError(cfg.transport)
    if comm.Get_rank() == cfg.master_rank:
        return run_master(cfg, comm)
    run_worker(cfg, comm)
    return None


def main(argv=None) -> int:
    os.environ.setdefault("TF_CPP_MIN_LOG_LEVEL", "3")  # harmless; mirrors main_manager.py:19
    cfg = parse_args(argv)
    # Under mpirun without mpi4py, or plain invocation: fall back to the built-in launcher.
    if cfg.transport == "mpi":
        from .parallel import mpi
        if not mpi.available():
            print("mpi4py not importable: degrading to the built-in local launcher")
            cfg = cfg.replace(transport="local")
    run(cfg)
    return 0
