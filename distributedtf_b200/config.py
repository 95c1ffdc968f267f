"""Typed run configuration + command line.

Parity: replaces the reference's two configuration systems: the
edit-the-source constants of ``main_manager.py:32-44`` (``master_rank,
train_round, population_size, epochs_per_round, do_exploit, do_explore,
target_model`` + the positional ``sys.argv[1]`` population override,
``:39-40``) and the absl-flags-as-global-dict of the CIFAR path
(``cifar10_main.py:239-297``).  Hyper-parameters travel explicitly with each
member; nothing here is global mutable state.
"""
from __future__ import annotations

import argparse
import dataclasses
from dataclasses import dataclass, field
from typing import List, Optional


@dataclass
class RunConfig:
    # --- reference constants (main_manager.py:32-44) ---
    master_rank: int = 0
    train_round: int = 20
    population_size: int = 20
    epochs_per_round: int = 1
    do_exploit: bool = True
    do_explore: bool = True
    target_model: str = "mnist"           # toy | mnist | cifar10
    # --- topology ---
    workers: int = 1                      # worker endpoints (reference: mpirun -n = workers + 1)
    transport: str = "local"              # local | tcp | gloo | mpi
    master_addr: str = "127.0.0.1"
    master_port: int = 29611
    dp_size: int = 1                      # GPUs per member (1 | 2 | 4)
    members_in_flight: int = 1            # members of one worker trained concurrently (own stream + host thread)
    sm_limit: int = 0                     # > 0: SMs a member's persistent kernels occupy (0 = all)
    checkpoint_every: int = 1             # rounds between member checkpoints (0 = last round only); 'file' exploit needs 1
    async_checkpoint: Optional[bool] = None   # None: on unless exploit_path == 'file' (native engines only)
    # --- data plane ---
    exploit_path: str = "file"            # file | p2p | nccl | host
    scoreboard: bool = False              # eval scores reach the master through the device score board
    engine: str = "auto"                  # native (sm_100a kernels) | eager (torch oracle) | auto
    dtype: str = "bf16"                   # bf16 | fp32 | fp16 (compute dtype of the native engine)
    # --- model / data ---
    resnet_size: int = 32
    resnet_version: int = 2
    synthetic: bool = True
    data_dir: str = "./datasets"
    max_train_steps: Optional[int] = None  # cap steps per epoch (synthetic / smoke runs)
    eval_batches: Optional[int] = None
    # --- housekeeping ---
    save_root: str = "./savedata"
    results_file: str = "test_results.txt"
    seed: Optional[int] = None
    recv_timeout: Optional[float] = None
    clean: bool = True                    # wipe save_root at start (main_manager.py:49-50)
    resume: bool = False                  # keep save_root, members restore their checkpoints
    initial_hp: Optional[str] = None      # JSON written by a previous run (initial_hp.json)
    inject_nan: List[str] = field(default_factory=list)   # "member_id@round" fault injection
    verbose: bool = True
    hooks: List[str] = field(default_factory=lambda: ["LoggingTensorHook"])
    benchmark_logger_type: str = "BaseBenchmarkLogger"
    benchmark_log_dir: Optional[str] = None

    def validate(self) -> "RunConfig":
        if self.workers < 1:
            raise ValueError("need at least one worker (the reference needs mpirun -n >= 2)")
        if self.checkpoint_every != 1 and self.exploit_path == "file" and self.do_exploit:
            raise ValueError("exploit_path='file' copies checkpoint files every round: checkpoint_every must be 1")
        if self.dp_size not in (1, 2, 4):
            raise ValueError("dp_size must be 1, 2 or 4")
        if self.target_model not in ("toy", "mnist", "cifar10"):
            raise ValueError("target_model must be toy | mnist | cifar10")
        if self.resnet_size % 6 != 2:
            raise ValueError("resnet_size must be 6n + 2")
        if self.resnet_version not in (1, 2):
            raise ValueError("resnet_version must be 1 or 2")
        if self.resnet_version == 1 and self.dtype == "fp16":
            # reference resnet_run_loop.py:542-549
            raise ValueError("ResNet v1 is not supported with fp16; use version 2")
        if self.resume:
            self.clean = False
        return self

    def replace(self, **kw) -> "RunConfig":
        return dataclasses.replace(self, **kw)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(
        prog="python -m distributedtf_b200",
        description="Population-based training on B200 (master/worker PBT).")
    p.add_argument("population", nargs="?", type=int, default=None,
                   help="population size (positional, like the reference's sys.argv[1])")
    p.add_argument("--model", dest="target_model", choices=["toy", "mnist", "cifar10"], default="mnist")
    p.add_argument("--population", dest="population_size", type=int, default=20)
    p.add_argument("--rounds", dest="train_round", type=int, default=20)
    p.add_argument("--epochs-per-round", type=int, default=1)
    p.add_argument("--exploit", dest="do_exploit", action=argparse.BooleanOptionalAction, default=True)
    p.add_argument("--explore", dest="do_explore", action=argparse.BooleanOptionalAction, default=True)
    p.add_argument("--workers", type=int, default=1)
    p.add_argument("--transport", choices=["local", "tcp", "torchrun", "gloo", "mpi"], default="local")
    p.add_argument("--master-addr", default="127.0.0.1")
    p.add_argument("--master-port", type=int, default=29611)
    p.add_argument("--dp-size", type=int, choices=[1, 2, 4], default=1)
    p.add_argument("--members-in-flight", type=int, default=1,
                   help="members of one worker trained at once (own stream + host thread each)")
    p.add_argument("--checkpoint-every", type=int, default=1,
                   help="rounds between member checkpoints (0 = only after the last round)")
    p.add_argument("--async-checkpoint", action=argparse.BooleanOptionalAction, default=None)
    p.add_argument("--sm-limit", type=int, default=0, help="SMs a member's persistent kernels occupy (0 = all)")
    p.add_argument("--exploit-path", choices=["file", "p2p", "nccl", "host"], default="file")
    p.add_argument("--scoreboard", action="store_true", help="rank by the device-side score board")
    p.add_argument("--engine", choices=["auto", "native", "eager"], default="auto")
    p.add_argument("--dtype", choices=["bf16", "fp32", "fp16"], default="bf16")
    p.add_argument("--resnet-size", type=int, default=32)
    p.add_argument("--resnet-version", type=int, choices=[1, 2], default=2)
    p.add_argument("--synthetic", action=argparse.BooleanOptionalAction, default=True)
    p.add_argument("--data-dir", default="./datasets")
    p.add_argument("--max-train-steps", type=int, default=None)
    p.add_argument("--eval-batches", type=int, default=None)
    p.add_argument("--save-root", default="./savedata")
    p.add_argument("--results-file", default="test_results.txt")
    p.add_argument("--seed", type=int, default=None)
    p.add_argument("--recv-timeout", type=float, default=None)
    p.add_argument("--resume", action="store_true")
    p.add_argument("--initial-hp", default=None)
    p.add_argument("--inject-nan", action="append", default=[], metavar="ID@ROUND")
    p.add_argument("--quiet", dest="verbose", action="store_false")
    p.add_argument("--hooks", nargs="*", default=["LoggingTensorHook"])
    p.add_argument("--benchmark-logger-type", default="BaseBenchmarkLogger")
    p.add_argument("--benchmark-log-dir", default=None)
    return p


def parse_args(argv=None) -> RunConfig:
    ns = build_parser().parse_args(argv)
    kw = vars(ns)
    positional = kw.pop("population")
    if positional is not None:
        kw["population_size"] = positional
    return RunConfig(**kw).validate()
