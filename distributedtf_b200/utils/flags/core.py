"""Flag groups for the model harness.

Parity: reference ``official/utils/flags/{core,_base,_performance,_misc,_benchmark,_device,
_conventions}.py``.  The reference defines absl flags as process-global mutable state and wipes +
re-defines them on every member call (``cifar10_main.py:325-329``).  Here a ``FlagSet`` is an
ordinary object: groups add typed definitions to it, ``parse`` returns an immutable namespace,
validators run at parse time, nothing is global.  Names, defaults and validators follow the
reference (``num_gpus`` default forced to 1 as in ``_base.py:102``; ``loss_scale`` must be > 0;
the file logger needs a log dir; TPU flags are accepted but inert).
"""
from __future__ import annotations

import argparse
import functools
import multiprocessing
import textwrap
from typing import Any, Callable, Dict, List, Optional, Sequence

import torch

help_wrap = functools.partial(textwrap.fill, width=80)

# dtype name -> (torch dtype, default loss scale)   (reference _performance.py:30-33)
DTYPE_MAP = {"fp16": (torch.float16, 128), "bf16": (torch.bfloat16, 1), "fp32": (torch.float32, 1),
             "tf32": (torch.float32, 1)}      # tf32: fp32 storage, tf32 tensor-core math (native engine)


class FlagSet:
    def __init__(self, description: str = ""):
        self._parser = argparse.ArgumentParser(description=description, allow_abbrev=False)
        self._validators: List[Callable[[argparse.Namespace], Optional[str]]] = []
        self._defined: Dict[str, Any] = {}
        self.key_flags: List[str] = []

    def define(self, name: str, short: Optional[str] = None, default: Any = None, help: str = "",
               type: Any = str, choices: Optional[Sequence[Any]] = None, nargs: Optional[str] = None) -> None:
        if name in self._defined:
            return
        # both spellings are accepted: --data_dir (the reference's absl names) and --data-dir (the PBT CLI)
        names = ["--" + name] + (["--" + name.replace("_", "-")] if "_" in name else []) + (["-" + short] if short else [])
        kw: Dict[str, Any] = dict(default=default, help=help_wrap(help))
        if type is bool:
            kw["action"] = argparse.BooleanOptionalAction
        else:
            kw["type"] = type
            if choices is not None:
                kw["choices"] = list(choices)
            if nargs:
                kw["nargs"] = nargs
        self._parser.add_argument(*names, dest=name, **kw)
        self._defined[name] = default
        self.key_flags.append(name)

    def validator(self, fn: Callable[[argparse.Namespace], Optional[str]]) -> None:
        self._validators.append(fn)

    def set_defaults(self, **kw) -> None:
        unknown = [k for k in kw if k not in self._defined]
        if unknown:
            raise KeyError(f"unknown flags: {unknown}")
        self._parser.set_defaults(**kw)
        self._defined.update(kw)

    @property
    def parser(self) -> argparse.ArgumentParser:
        """The underlying parser, for callers that add their own (non-harness) arguments next to the groups."""
        return self._parser

    def validate(self, ns: argparse.Namespace) -> argparse.Namespace:
        for v in self._validators:
            msg = v(ns)
            if msg:
                raise ValueError(msg)
        return ns

    def parse(self, argv: Optional[Sequence[str]] = None) -> argparse.Namespace:
        return self.validate(self._parser.parse_args(list(argv) if argv is not None else []))


def define_base(fs: FlagSet, data_dir=True, model_dir=True, clean=True, train_epochs=True,
                epochs_between_evals=True, stop_threshold=True, batch_size=True, num_gpu=True, hooks=True,
                export_dir=True) -> List[str]:
    before = len(fs.key_flags)
    if data_dir:
        fs.define("data_dir", "dd", "/tmp", "The location of the input data.")
    if model_dir:
        fs.define("model_dir", "md", "/tmp", "The location of the model checkpoint files.")
    if clean:
        fs.define("clean", None, False, "If set, model_dir will be removed if it exists.", type=bool)
    if train_epochs:
        fs.define("train_epochs", "te", 1, "The number of epochs used to train.", type=int)
    if epochs_between_evals:
        fs.define("epochs_between_evals", "ebe", 1, "The number of training epochs to run between evaluations.", type=int)
    if stop_threshold:
        fs.define("stop_threshold", "st", None, "If passed, training stops at the earlier of train_epochs and when "
                  "the evaluation metric is greater than or equal to stop_threshold.", type=float)
    if batch_size:
        fs.define("batch_size", "bs", 32, "Global batch size; it is split evenly over the GPUs of a "
                  "data-parallel member.", type=int)
    if num_gpu:
        fs.define("num_gpus", "ng", 1, "How many GPUs one member uses (1, 2 or 4).", type=int)
    if hooks:
        fs.define("hooks", "hk", ["LoggingTensorHook"], "Names of train hooks: LoggingTensorHook, ProfilerHook, "
                  "ExamplesPerSecondHook, LoggingMetricHook (case insensitive).", nargs="*")
    if export_dir:
        fs.define("export_dir", "ed", None, "If set, an inference bundle is exported to this directory at the end.")
    return fs.key_flags[before:]


def define_performance(fs: FlagSet, num_parallel_calls=True, inter_op=True, intra_op=True,
                       synthetic_data=True, max_train_steps=True, dtype=True, all_reduce_alg=True) -> List[str]:
    before = len(fs.key_flags)
    if num_parallel_calls:
        fs.define("num_parallel_calls", "npc", multiprocessing.cpu_count(), "Host threads of the input pipeline.", type=int)
    if inter_op:
        fs.define("inter_op_parallelism_threads", "inter", 0, "Kept for compatibility; streams replace it.", type=int)
    if intra_op:
        fs.define("intra_op_parallelism_threads", "intra", 0, "Kept for compatibility.", type=int)
    if synthetic_data:
        fs.define("use_synthetic_data", "synth", False, "Use generated data of the dataset's shape.", type=bool)
    if max_train_steps:
        fs.define("max_train_steps", "mts", None, "Cap on training steps per cycle (tests / profiling).", type=int)
    if dtype:
        fs.define("dtype", "dt", "fp32", "Compute dtype.", choices=list(DTYPE_MAP))
        fs.define("loss_scale", "ls", None, "Static loss scale; default depends on dtype (128 for fp16).", type=int)
        fs.validator(lambda ns: None if ns.loss_scale is None or ns.loss_scale > 0 else "loss_scale should be a positive integer.")
    if all_reduce_alg:
        fs.define("all_reduce_alg", "ara", None, "Gradient all-reduce of a data-parallel member: fused_p2p "
                  "(one-shot kernel over peer memory, default) or nccl (library baseline).")
    return fs.key_flags[before:]


def define_image(fs: FlagSet, data_format=True) -> List[str]:
    before = len(fs.key_flags)
    if data_format:
        fs.define("data_format", "df", "channels_last", "The native engine is NHWC end to end; channels_first is "
                  "accepted by the eager engine only.", choices=["channels_first", "channels_last"])
    return fs.key_flags[before:]


def define_benchmark(fs: FlagSet, benchmark_log_dir=True, bigquery_uploader=True) -> List[str]:
    before = len(fs.key_flags)
    fs.define("benchmark_logger_type", "blt", "BaseBenchmarkLogger", "Where benchmark metrics go.",
              choices=["BaseBenchmarkLogger", "BenchmarkFileLogger", "BenchmarkBigQueryLogger"])
    fs.define("benchmark_test_id", "bti", None, "Test id recorded with the run info.")
    if benchmark_log_dir:
        fs.define("benchmark_log_dir", "bld", None, "Directory of metric.log / benchmark_run.log.")
    if bigquery_uploader:
        for n, d in (("gcp_project", None), ("bigquery_data_set", "test_benchmark"), ("bigquery_run_table", "benchmark_run"),
                     ("bigquery_run_status_table", "benchmark_run_status"), ("bigquery_metric_table", "benchmark_metric")):
            fs.define(n, None, d, "BigQuery upload target (no network on this platform: accepted, unused).")
    fs.validator(lambda ns: "--benchmark_logger_type=BenchmarkFileLogger will require --benchmark_log_dir being set"
                 if ns.benchmark_logger_type == "BenchmarkFileLogger" and not getattr(ns, "benchmark_log_dir", None) else None)
    return fs.key_flags[before:]


def define_device(fs: FlagSet, tpu=True) -> List[str]:
    before = len(fs.key_flags)
    if tpu:
        for n in ("tpu", "tpu_zone", "tpu_gcp_project"):
            fs.define(n, None, None, "TPU flags of the reference: accepted, inert (no TPU path).")
        fs.define("num_tpu_shards", None, 8, "inert", type=int)
    return fs.key_flags[before:]


def get_num_gpus(ns) -> int:
    if ns.num_gpus != -1:
        return ns.num_gpus
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def get_torch_dtype(ns) -> torch.dtype:
    return DTYPE_MAP[ns.dtype][0]


get_tf_dtype = get_torch_dtype     # reference name


def get_loss_scale(ns) -> int:
    return ns.loss_scale if ns.loss_scale is not None else DTYPE_MAP[ns.dtype][1]


def define_resnet_flags(fs: Optional[FlagSet] = None, resnet_size_choices: Optional[Sequence[str]] = None) -> FlagSet:
    """Reference ``resnet_run_loop.define_resnet_flags`` (:519-549)."""
    fs = fs or FlagSet("ResNet")
    define_base(fs)
    define_performance(fs, num_parallel_calls=False)
    define_image(fs)
    define_benchmark(fs)
    fs.define("resnet_version", "rv", "2", "Version of ResNet (1 or 2).", choices=["1", "2"])
    fs.define("resnet_size", "rs", "50", "The size of the ResNet model to use.",
              choices=list(resnet_size_choices) if resnet_size_choices else None)
    fs.validator(lambda ns: "ResNet version 1 is not currently supported with fp16. Please use version 2 instead."
                 if ns.dtype == "fp16" and str(ns.resnet_version) == "1" else None)
    return fs
