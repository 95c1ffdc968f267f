"""Training hooks.

Parity: reference ``official/utils/logs/{hooks,metric_hook,hooks_helper}.py``.  TensorFlow's
``SessionRunHook`` protocol (begin / before_run / after_run / end) becomes a plain per-step protocol:
``begin()``, ``after_step(step, values)`` with ``values`` a dict of host scalars *or zero-dim device
tensors* (converted only when a hook actually logs, so hooks never force a sync on other steps),
``end()``.  Registry names match the reference: ``loggingtensorhook``, ``profilerhook``,
``examplespersecondhook``, ``loggingmetrichook``.
"""
from __future__ import annotations

import json
import logging
import os
import time
from typing import Any, Callable, Dict, Iterable, List, Optional

from . import logger as _logger

_log = logging.getLogger("distributedtf_b200.hooks")
_TENSORS_TO_LOG = {x: x for x in ["learning_rate", "cross_entropy", "train_accuracy"]}


def _scalar(v: Any) -> float:
    return float(v.item()) if hasattr(v, "item") else float(v)


class Hook:
    def begin(self) -> None: ...
    def after_step(self, step: int, values: Dict[str, Any]) -> None: ...
    def end(self) -> None: ...


class _EveryN:
    def __init__(self, every_n_steps: Optional[int], every_n_secs: Optional[float]):
        if (every_n_steps is None) == (every_n_secs is None):
            raise ValueError("exactly one of every_n_steps and every_n_secs should be provided.")
        self.every_n_steps, self.every_n_secs = every_n_steps, every_n_secs
        self.last_step: Optional[int] = None
        self.last_time: Optional[float] = None

    def should_trigger(self, step: int) -> bool:
        if self.last_step is None:
            return True
        if self.every_n_steps is not None:
            return step >= self.last_step + self.every_n_steps
        return time.time() >= self.last_time + self.every_n_secs

    def update(self, step: int):
        now = time.time()
        elapsed = (None, None) if self.last_step is None else (now - self.last_time, step - self.last_step)
        self.last_step, self.last_time = step, now
        return elapsed


class ExamplesPerSecondHook(Hook):
    """Average and current examples/sec every N steps or seconds, after ``warm_steps``."""

    def __init__(self, batch_size: int, every_n_steps: Optional[int] = None, every_n_secs: Optional[float] = None,
                 warm_steps: int = 0, metric_logger=None):
        self._timer = _EveryN(every_n_steps, every_n_secs)
        self._batch_size, self._warm_steps = batch_size, warm_steps
        self._logger = metric_logger or _logger.BaseBenchmarkLogger()
        self._step_train_time, self._total_steps = 0.0, 0

    def after_step(self, step: int, values: Dict[str, Any]) -> None:
        if step < self._warm_steps or not self._timer.should_trigger(step):
            return
        elapsed_time, elapsed_steps = self._timer.update(step)
        if elapsed_time is None:
            return
        self._step_train_time += elapsed_time
        self._total_steps += elapsed_steps
        average = self._batch_size * (self._total_steps / self._step_train_time)
        current = self._batch_size * (elapsed_steps / elapsed_time)
        self._logger.log_metric("average_examples_per_sec", average, global_step=step)
        self._logger.log_metric("current_examples_per_sec", current, global_step=step)


class LoggingTensorHook(Hook):
    def __init__(self, tensors: Iterable[str], every_n_iter: int = 100, sink: Optional[Callable[[str], None]] = None):
        self._names = list(tensors)
        self._timer = _EveryN(every_n_iter, None)
        self._sink = sink or _log.info

    def after_step(self, step: int, values: Dict[str, Any]) -> None:
        if self._timer.should_trigger(step):
            self._timer.update(step)
            self._sink(", ".join(f"{k} = {_scalar(values[k]):.6g}" for k in self._names if k in values) + f" (step {step})")


class LoggingMetricHook(Hook):
    """Sends the named values to a benchmark logger instead of the text log."""

    def __init__(self, tensors: Iterable[str], metric_logger=None, every_n_iter: Optional[int] = None,
                 every_n_secs: Optional[float] = None, at_end: bool = False):
        if metric_logger is None:
            raise ValueError("metric_logger should be provided.")
        self._names, self._logger, self._at_end = list(tensors), metric_logger, at_end
        self._timer = _EveryN(every_n_iter, every_n_secs) if (every_n_iter or every_n_secs) else None
        self._last: Dict[str, Any] = {}
        self._last_step = 0

    def after_step(self, step: int, values: Dict[str, Any]) -> None:
        self._last, self._last_step = values, step
        if self._timer is not None and self._timer.should_trigger(step):
            self._timer.update(step)
            self._emit(step, values)

    def end(self) -> None:
        if self._at_end and self._last:
            self._emit(self._last_step, self._last)

    def _emit(self, step: int, values: Dict[str, Any]) -> None:
        for k in self._names:
            if k in values:
                self._logger.log_metric(k, _scalar(values[k]), global_step=step)


class ProfilerHook(Hook):
    """Chrome-trace timeline (one complete event per step) written to ``model_dir/timeline-<step>.json``
    every ``save_steps`` — the role of ``tf.train.ProfilerHook`` in the reference.  On a CUDA device the step
    durations are DEVICE times: one CUDA event is recorded on the member's stream after every step and the gaps
    between consecutive events are read when the trace is written (no synchronisation in between); on the CPU they
    are host times."""

    def __init__(self, model_dir: str, save_steps: int = 1000):
        self._dir, self._save_steps = model_dir, save_steps
        self._events: List[Dict[str, Any]] = []
        self._t_prev: Optional[float] = None
        self._cuda_events: List[Any] = []          # (step, host time, cuda event)

    @staticmethod
    def _cuda_live() -> bool:
        try:
            import torch
            return torch.cuda.is_available() and torch.cuda.is_initialized()
        except Exception:
            return False

    def after_step(self, step: int, values: Dict[str, Any]) -> None:
        now = time.time()
        if self._cuda_live():
            import torch
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._cuda_events.append((step, now, ev))
        elif self._t_prev is not None:
            self._events.append({"name": "train_step", "ph": "X", "pid": os.getpid(), "tid": 0,
                                 "ts": self._t_prev * 1e6, "dur": (now - self._t_prev) * 1e6, "args": {"step": step}})
        self._t_prev = now
        if step % self._save_steps == 0:
            self._flush(step)

    def end(self) -> None:
        if self._cuda_events or self._events:
            self._flush(self._cuda_events[-1][0] if self._cuda_events else 0)

    def _flush(self, step: int) -> None:
        if len(self._cuda_events) > 1:
            self._cuda_events[-1][2].synchronize()
            for (s0, t0, e0), (s1, _, e1) in zip(self._cuda_events[:-1], self._cuda_events[1:]):
                self._events.append({"name": "train_step", "ph": "X", "pid": os.getpid(), "tid": 0, "ts": t0 * 1e6,
                                     "dur": e0.elapsed_time(e1) * 1e3, "args": {"step": s1, "clock": "device"}})
            self._cuda_events = self._cuda_events[-1:]
        if not self._events:
            return
        os.makedirs(self._dir, exist_ok=True)
        with open(os.path.join(self._dir, f"timeline-{step}.json"), "w") as f:
            json.dump({"traceEvents": self._events}, f)
        self._events = []


def get_logging_tensor_hook(every_n_iter=100, tensors_to_log=None, **kwargs):
    return LoggingTensorHook(tensors_to_log or _TENSORS_TO_LOG, every_n_iter=every_n_iter)


def get_profiler_hook(model_dir, save_steps=1000, **kwargs):
    return ProfilerHook(model_dir, save_steps=save_steps)


def get_examples_per_second_hook(every_n_steps=100, batch_size=128, warm_steps=5, **kwargs):
    return ExamplesPerSecondHook(batch_size, every_n_steps=every_n_steps, warm_steps=warm_steps,
                                 metric_logger=_logger.get_benchmark_logger())


def get_logging_metric_hook(tensors_to_log=None, every_n_secs=600, **kwargs):
    return LoggingMetricHook(tensors_to_log or _TENSORS_TO_LOG, metric_logger=_logger.get_benchmark_logger(),
                             every_n_secs=every_n_secs)


HOOKS = {"loggingtensorhook": get_logging_tensor_hook, "profilerhook": get_profiler_hook,
         "examplespersecondhook": get_examples_per_second_hook, "loggingmetrichook": get_logging_metric_hook}


def get_train_hooks(name_list, use_tpu=False, **kwargs) -> List[Hook]:
    """Reference ``hooks_helper.get_train_hooks``: build hooks by (case-insensitive) name."""
    if not name_list:
        return []
    if use_tpu:
        _log.warning("hooks are not TPU compatible; there is no TPU path — ignoring use_tpu")
    out = []
    for name in name_list:
        factory = HOOKS.get(name.strip().lower())
        if factory is None:
            raise ValueError("Unrecognized training hook requested: {}".format(name))
        out.append(factory(**kwargs))
    return out
