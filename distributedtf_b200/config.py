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
    scoreboard: Optional[bool] = None     # eval scores reach the master through the device score board
                                          # (None: on for exploit_path == "p2p", off otherwise)
    engine: str = "auto"                  # native (sm_100a kernels) | eager (torch oracle) | auto
    dtype: str = "bf16"                   # bf16 | tf32 | fp32 | fp16 (compute dtype of the native engine; fp32 = tf32 MMA)
    loss_scale: Optional[int] = None      # static loss scale; None = 128 for fp16, 1 otherwise (flags.get_loss_scale)
    all_reduce_alg: Optional[str] = None  # data-parallel member: fused_p2p (default) | nccl (library baseline)
    # --- model / data ---
    resnet_size: int = 32
    resnet_version: int = 2
    resnet_bottleneck: bool = False       # bottleneck blocks (resnet_model.py:215-320); resnet_size is then 9n + 2
    synthetic: bool = True
    data_dir: str = "./datasets"
    max_train_steps: Optional[int] = None  # cap steps per epoch (synthetic / smoke runs)
    eval_batches: Optional[int] = None
    # --- run-loop controls of the model harness (resnet_run_loop.py:434-448,505-514) ---
    epochs_between_evals: int = 1         # training epochs between two evaluations inside one train() call
    stop_threshold: Optional[float] = None   # a member stops training early once its eval accuracy reaches this
    export_dir: Optional[str] = None      # inference bundle of every member, written at the end of training
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
        if self.resnet_size % (9 if self.resnet_bottleneck else 6) != 2:
            raise ValueError("resnet_size must be 6n + 2 (9n + 2 with bottleneck blocks)")
        if self.resnet_version not in (1, 2):
            raise ValueError("resnet_version must be 1 or 2")
        if self.resnet_version == 1 and self.dtype == "fp16":
            # reference resnet_run_loop.py:542-549
            raise ValueError("ResNet v1 is not supported with fp16; use version 2")
        if self.dtype not in ("bf16", "tf32", "fp32", "fp16"):
            raise ValueError("dtype must be bf16 | tf32 | fp32 | fp16")
        if self.loss_scale is not None and self.loss_scale <= 0:
            raise ValueError("loss_scale should be a positive integer.")
        if self.epochs_between_evals < 1:
            raise ValueError("epochs_between_evals must be >= 1")
        # strategy selection of the reference (distribution_utils.py:24-56): validates dp_size / all_reduce_alg
        from .utils.misc.distribution_utils import get_distribution_strategy
        get_distribution_strategy(self.dp_size, self.all_reduce_alg)
        if self.scoreboard is None:
            self.scoreboard = self.exploit_path == "p2p" and self.target_model != "toy"
        if self.resume:
            self.clean = False
        return self

    def replace(self, **kw) -> "RunConfig":
        return dataclasses.replace(self, **kw)


def harness_flags():
    """The model-harness flag groups of the reference (``official/utils/flags``: base, performance, benchmark)
    as one ``FlagSet``; the PBT-level arguments are added to the same parser by ``build_parser``.  Defaults
    differ from the reference where the PBT runner differs: synthetic data and bf16 are on by default."""
    from .utils.flags import core as flags_core
    fs = flags_core.FlagSet("Population-based training on B200 (master/worker PBT).")
    fs.parser.prog = "python -m distributedtf_b200"
    flags_core.define_base(fs, model_dir=False, clean=False, train_epochs=False, batch_size=False, num_gpu=False)
    flags_core.define_performance(fs, num_parallel_calls=False, inter_op=False, intra_op=False, synthetic_data=False)
    flags_core.define_benchmark(fs, bigquery_uploader=False)
    fs.set_defaults(data_dir="./datasets", dtype="bf16")
    return fs


def build_parser() -> argparse.ArgumentParser:
    fs = harness_flags()
    p = fs.parser
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
    p.add_argument("--dp-size", "--num_gpus", "--num-gpus", dest="dp_size", type=int, choices=[1, 2, 4], default=1,
                   help="GPUs per member (the reference's num_gpus, which it forces to 1)")
    p.add_argument("--members-in-flight", type=int, default=1,
                   help="members of one worker trained at once (own stream + host thread each)")
    p.add_argument("--checkpoint-every", type=int, default=1,
                   help="rounds between member checkpoints (0 = only after the last round)")
    p.add_argument("--async-checkpoint", action=argparse.BooleanOptionalAction, default=None)
    p.add_argument("--sm-limit", type=int, default=0, help="SMs a member's persistent kernels occupy (0 = all)")
    p.add_argument("--exploit-path", choices=["file", "p2p", "nccl", "host"], default="file")
    p.add_argument("--scoreboard", action=argparse.BooleanOptionalAction, default=None,
                   help="rank by the device-side score board (default: on for the device exploit paths)")
    p.add_argument("--engine", choices=["auto", "native", "eager"], default="auto")
    p.add_argument("--resnet-size", "--resnet_size", dest="resnet_size", type=int, default=32)
    p.add_argument("--resnet-version", "--resnet_version", dest="resnet_version", type=int, choices=[1, 2], default=2)
    p.add_argument("--resnet-bottleneck", "--resnet_bottleneck", dest="resnet_bottleneck",
                   action=argparse.BooleanOptionalAction, default=False,
                   help="build the three stages from bottleneck blocks (resnet_size = 9n + 2)")
    p.add_argument("--synthetic", "--use_synthetic_data", dest="synthetic", action=argparse.BooleanOptionalAction,
                   default=True)
    p.add_argument("--eval-batches", type=int, default=None)
    p.add_argument("--save-root", default="./savedata")
    p.add_argument("--results-file", default="test_results.txt")
    p.add_argument("--seed", type=int, default=None)
    p.add_argument("--recv-timeout", type=float, default=None)
    p.add_argument("--resume", action="store_true")
    p.add_argument("--initial-hp", default=None)
    p.add_argument("--inject-nan", action="append", default=[], metavar="ID@ROUND")
    p.add_argument("--quiet", dest="verbose", action="store_false")
    p._dtf_flagset = fs
    return p


def parse_args(argv=None) -> RunConfig:
    parser = build_parser()
    ns = parser._dtf_flagset.validate(parser.parse_args(argv))    # group validators: loss_scale > 0, file logger dir
    kw = vars(ns)
    positional = kw.pop("population")
    if positional is not None:
        kw["population_size"] = positional
    kw.pop("benchmark_test_id", None)
    return RunConfig(**kw).validate()
