"""Benchmark loggers.

Parity: reference ``official/utils/logs/logger.py`` — a process-wide logger behind a lock
(``config_benchmark_logger`` / ``get_benchmark_logger`` / ``benchmark_context``), three sinks
(``BaseBenchmarkLogger`` -> python logging, ``BenchmarkFileLogger`` -> ``metric.log`` +
``benchmark_run.log`` JSON lines, ``BenchmarkBigQueryLogger`` -> threaded upload through a pluggable uploader; offline spool here), the metric
schema ``{name, value, unit, global_step, timestamp, extras}`` and the run-info collectors
(framework version, ``DTF_*``/``CUDA_*`` environment, run parameters, CPU, memory, GPUs, test
environment).
"""
from __future__ import annotations

import contextlib
import datetime
import json
import logging
import multiprocessing
import numbers
import os
import threading
import uuid
from typing import Any, Dict, List, Optional

from . import cloud_lib

METRIC_LOG_FILE_NAME = "metric.log"
BENCHMARK_RUN_LOG_FILE_NAME = "benchmark_run.log"
_DATE_TIME_FORMAT_PATTERN = "%Y-%m-%dT%H:%M:%S.%fZ"
GCP_TEST_ENV = "GCP"
RUN_STATUS_SUCCESS, RUN_STATUS_FAILURE, RUN_STATUS_RUNNING = "success", "failure", "running"

_log = logging.getLogger("distributedtf_b200.benchmark")
_benchmark_logger = None
_logger_lock = threading.Lock()


def config_benchmark_logger(flag_obj: Any = None):
    global _benchmark_logger
    with _logger_lock:
        kind = getattr(flag_obj, "benchmark_logger_type", "BaseBenchmarkLogger") if flag_obj is not None \
            else "BaseBenchmarkLogger"
        if kind == "BaseBenchmarkLogger":
            _benchmark_logger = BaseBenchmarkLogger()
        elif kind == "BenchmarkFileLogger":
            _benchmark_logger = BenchmarkFileLogger(flag_obj.benchmark_log_dir)
        elif kind == "BenchmarkBigQueryLogger":
            uploader = make_bigquery_uploader(getattr(flag_obj, "gcp_project", None),
                                              getattr(flag_obj, "benchmark_log_dir", None))
            _benchmark_logger = BenchmarkBigQueryLogger(
                uploader, getattr(flag_obj, "bigquery_data_set", "test_benchmark"),
                getattr(flag_obj, "bigquery_run_table", "benchmark_run"),
                getattr(flag_obj, "bigquery_run_status_table", "benchmark_run_status"),
                getattr(flag_obj, "bigquery_metric_table", "benchmark_metric"), run_id=str(uuid.uuid4()))
        else:
            raise ValueError("Unrecognized benchmark_logger_type: %s" % kind)
        return _benchmark_logger


def get_benchmark_logger():
    if _benchmark_logger is None:
        config_benchmark_logger()
    return _benchmark_logger


@contextlib.contextmanager
def benchmark_context(flag_obj: Any):
    logger = config_benchmark_logger(flag_obj)
    try:
        yield
        logger.on_finish(RUN_STATUS_SUCCESS)
    except Exception:
        logger.on_finish(RUN_STATUS_FAILURE)
        raise


class BaseBenchmarkLogger:
    def log_evaluation_result(self, eval_results: Dict[str, Any]) -> None:
        if not isinstance(eval_results, dict):
            _log.warning("eval_results should be dictionary for logging. Got %s", type(eval_results))
            return
        global_step = eval_results.get("global_step")
        for key in sorted(eval_results):
            if key != "global_step":
                self.log_metric(key, eval_results[key], global_step=global_step)

    def log_metric(self, name, value, unit=None, global_step=None, extras=None) -> None:
        metric = _process_metric_to_json(name, value, unit, global_step, extras)
        if metric:
            _log.info("Benchmark metric: %s", metric)

    def log_run_info(self, model_name, dataset_name, run_params, test_id=None) -> None:
        _log.info("Benchmark run: %s", _gather_run_info(model_name, dataset_name, run_params, test_id))

    def on_finish(self, status) -> None:
        pass


class BenchmarkFileLogger(BaseBenchmarkLogger):
    def __init__(self, logging_dir: str):
        self._logging_dir = logging_dir
        os.makedirs(logging_dir, exist_ok=True)
        self._metric_file = open(os.path.join(logging_dir, METRIC_LOG_FILE_NAME), "a")

    def log_metric(self, name, value, unit=None, global_step=None, extras=None) -> None:
        metric = _process_metric_to_json(name, value, unit, global_step, extras)
        if metric:
            try:
                json.dump(metric, self._metric_file)
                self._metric_file.write("\n")
                self._metric_file.flush()
            except (TypeError, ValueError) as e:
                _log.warning("Failed to dump metric to log file: name %s, value %s, error %s", name, value, e)

    def log_run_info(self, model_name, dataset_name, run_params, test_id=None) -> None:
        info = _gather_run_info(model_name, dataset_name, run_params, test_id)
        with open(os.path.join(self._logging_dir, BENCHMARK_RUN_LOG_FILE_NAME), "w") as f:
            try:
                json.dump(info, f)
                f.write("\n")
            except (TypeError, ValueError) as e:
                _log.warning("Failed to dump benchmark run info to log file: %s", e)

    def on_finish(self, status) -> None:
        self._metric_file.flush()
        self._metric_file.close()


class SpoolUploader:
    """Stand-in for ``benchmark_uploader.BigQueryUploader`` on machines without network /
    ``google-cloud-bigquery``: rows that would be streamed to ``<dataset>.<table>`` are appended to
    ``<dir>/bigquery_spool/<dataset>.<table>.jsonl`` for a later upload.  Same three calls as the reference's
    uploader (``logger.py:221-300`` uses ``upload_benchmark_run_json / upload_metric_json /
    insert_run_status / update_run_status``)."""

    def __init__(self, spool_dir: str):
        self.dir = os.path.join(spool_dir, "bigquery_spool")
        os.makedirs(self.dir, exist_ok=True)
        self._lock = threading.Lock()

    def _append(self, dataset, table, row) -> None:
        with self._lock, open(os.path.join(self.dir, f"{dataset}.{table}.jsonl"), "a") as f:
            f.write(json.dumps(row, default=str) + "\n")

    def upload_benchmark_run_json(self, dataset, table, run_id, run_json) -> None:
        self._append(dataset, table, dict(run_json, model_id=run_id))

    def upload_benchmark_metric_json(self, dataset, table, run_id, metric_json_list) -> None:
        for m in metric_json_list:
            self._append(dataset, table, dict(m, run_id=run_id))

    def insert_run_status(self, dataset, table, run_id, run_status) -> None:
        self._append(dataset, table, {"run_id": run_id, "status": run_status})

    def update_run_status(self, dataset, table, run_id, run_status) -> None:
        self._append(dataset, table, {"run_id": run_id, "status": run_status, "update": True})


def make_bigquery_uploader(gcp_project: Optional[str], spool_dir: Optional[str]):
    """The real uploader when ``google.cloud.bigquery`` imports and a project is named, else the spool."""
    if gcp_project:
        try:
            from google.cloud import bigquery  # type: ignore  # noqa: F401
            raise ImportError("streaming upload is not wired on this platform")   # no egress here
        except ImportError:
            pass
    return SpoolUploader(spool_dir or "/tmp/dtf_benchmark")


class BenchmarkBigQueryLogger(BaseBenchmarkLogger):
    """Uploads run info, metrics and the final status from background threads so that logging never blocks
    the training step (reference ``logs/logger.py:221-300``: one ``thread.start_new_thread`` per call)."""

    def __init__(self, bigquery_uploader, bigquery_data_set, bigquery_run_table, bigquery_run_status_table,
                 bigquery_metric_table, run_id):
        self._uploader = bigquery_uploader
        self._data_set = bigquery_data_set
        self._run_table = bigquery_run_table
        self._run_status_table = bigquery_run_status_table
        self._metric_table = bigquery_metric_table
        self._run_id = run_id
        self._threads: List[threading.Thread] = []

    def _spawn(self, fn, *args) -> None:
        t = threading.Thread(target=fn, args=args, daemon=True)
        t.start()
        self._threads.append(t)

    def log_metric(self, name, value, unit=None, global_step=None, extras=None) -> None:
        metric = _process_metric_to_json(name, value, unit, global_step, extras)
        if metric:
            self._spawn(self._uploader.upload_benchmark_metric_json, self._data_set, self._metric_table,
                        self._run_id, [metric])

    def log_run_info(self, model_name, dataset_name, run_params, test_id=None) -> None:
        info = _gather_run_info(model_name, dataset_name, run_params, test_id)
        self._spawn(self._uploader.upload_benchmark_run_json, self._data_set, self._run_table, self._run_id, info)
        self._spawn(self._uploader.insert_run_status, self._data_set, self._run_status_table, self._run_id,
                    RUN_STATUS_RUNNING)

    def on_finish(self, status) -> None:
        for t in self._threads:              # ordered after every earlier row of this run
            t.join()
        self._threads.clear()
        self._uploader.update_run_status(self._data_set, self._run_status_table, self._run_id, status)


def _gather_run_info(model_name, dataset_name, run_params, test_id) -> Dict[str, Any]:
    run_info: Dict[str, Any] = {
        "model_name": model_name, "dataset": {"name": dataset_name}, "machine_config": {}, "test_id": test_id,
        "run_date": datetime.datetime.utcnow().strftime(_DATE_TIME_FORMAT_PATTERN)}
    _collect_framework_info(run_info)
    _collect_environment_variables(run_info)
    _collect_run_params(run_info, run_params)
    _collect_cpu_info(run_info)
    _collect_gpu_info(run_info)
    _collect_memory_info(run_info)
    _collect_test_environment(run_info)
    return run_info


def _process_metric_to_json(name, value, unit=None, global_step=None, extras=None) -> Optional[Dict[str, Any]]:
    if not isinstance(value, numbers.Number):
        _log.warning("Metric value to log should be a number. Got %s", type(value))
        return None
    return {"name": name, "value": float(value), "unit": unit, "global_step": global_step,
            "timestamp": datetime.datetime.utcnow().strftime(_DATE_TIME_FORMAT_PATTERN),
            "extras": _convert_to_json_dict(extras)}


def _collect_framework_info(run_info) -> None:
    import torch
    from ... import __version__
    run_info["framework_version"] = {"version": __version__, "torch": torch.__version__,
                                     "cuda": torch.version.cuda}


def _collect_run_params(run_info, run_params) -> None:
    def process(name, value):
        type_check = {str: {"name": name, "string_value": value}, int: {"name": name, "long_value": value},
                      bool: {"name": name, "bool_value": str(value)}, float: {"name": name, "float_value": value}}
        return type_check.get(type(value), {"name": name, "string_value": str(value)})
    if run_params:
        run_info["run_parameters"] = [process(k, v) for k, v in sorted(run_params.items())]


def _collect_environment_variables(run_info) -> None:
    run_info["environment_variables"] = [{"name": k, "value": v} for k, v in sorted(os.environ.items())
                                         if k.startswith(("DTF_", "CUDA_", "NCCL_"))]


def _collect_cpu_info(run_info) -> None:
    cpu = {"num_cores": multiprocessing.cpu_count()}
    try:
        import cpuinfo
        info = cpuinfo.get_cpu_info()
        cpu["cpu_info"] = info.get("brand_raw", info.get("brand"))
        cpu["mhz_per_cpu"] = info.get("hz_advertised_friendly")
    except Exception:
        pass
    run_info["machine_config"]["cpu_info"] = cpu


def _collect_gpu_info(run_info) -> None:
    try:
        import torch
        if torch.cuda.is_available():
            run_info["machine_config"]["gpu_info"] = {"count": torch.cuda.device_count(),
                                                     "model": torch.cuda.get_device_name(0)}
    except Exception:
        pass


def _collect_memory_info(run_info) -> None:
    try:
        import psutil
        vmem = psutil.virtual_memory()
        run_info["machine_config"]["memory_total"] = vmem.total
        run_info["machine_config"]["memory_available"] = vmem.available
    except Exception:
        pass


def _collect_test_environment(run_info) -> None:
    if cloud_lib.on_gcp():
        run_info["test_environment"] = GCP_TEST_ENV


def _convert_to_json_dict(input_dict):
    if input_dict:
        return [{"name": k, "value": v} for k, v in sorted(input_dict.items())]
    return []
