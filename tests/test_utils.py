"""Utilities: flags, loggers, hooks, distribution strategy, model helpers, export, golden-data harness,
file_io, embedding_matmul.  Mirrors what the reference's official/utils/**/*_test.py cover."""
import json
import os
import time

import numpy as np
import pytest
import torch

from distributedtf_b200.utils.flags import core as flags_core
from distributedtf_b200.utils.logs import cloud_lib, hooks, logger
from distributedtf_b200.utils.misc import distribution_utils, model_helpers
from distributedtf_b200.utils.testing import integration, mock_lib, reference_data


# ------------------------------------------------------------------ flags (flags_test.py:38-96)
def _fs():
    fs = flags_core.FlagSet()
    flags_core.define_base(fs)
    flags_core.define_performance(fs)
    flags_core.define_image(fs)
    flags_core.define_benchmark(fs)
    return fs


def test_flag_defaults_and_overrides():
    fs = _fs()
    ns = fs.parse([])
    assert (ns.data_dir, ns.model_dir, ns.train_epochs, ns.epochs_between_evals, ns.batch_size, ns.num_gpus) == \
        ("/tmp", "/tmp", 1, 1, 32, 1)
    assert ns.hooks == ["LoggingTensorHook"] and ns.dtype == "fp32" and ns.data_format == "channels_last"
    fs.set_defaults(data_dir="dfgasf", model_dir="dfsdkjgbs", train_epochs=534, batch_size=256)
    ns = fs.parse([])
    assert (ns.data_dir, ns.model_dir, ns.train_epochs, ns.batch_size) == ("dfgasf", "dfsdkjgbs", 534, 256)
    with pytest.raises(KeyError):
        fs.set_defaults(not_a_flag=1)
    ns = fs.parse(["--batch_size", "99", "-te", "7", "--use_synthetic_data", "--no-clean"])
    assert ns.batch_size == 99 and ns.train_epochs == 7 and ns.use_synthetic_data is True and ns.clean is False


def test_dtype_and_loss_scale():
    fs = _fs()
    for dtype, want_t, want_ls in [("fp16", torch.float16, 128), ("fp32", torch.float32, 1), ("bf16", torch.bfloat16, 1)]:
        ns = fs.parse(["--dtype", dtype])
        assert flags_core.get_torch_dtype(ns) is want_t and flags_core.get_loss_scale(ns) == want_ls
        ns = fs.parse(["--dtype", dtype, "--loss_scale", "5"])
        assert flags_core.get_loss_scale(ns) == 5
    with pytest.raises(SystemExit):
        fs.parse(["--dtype", "int8"])
    with pytest.raises(ValueError):
        fs.parse(["--loss_scale", "0"])
    with pytest.raises(ValueError):
        fs.parse(["--benchmark_logger_type", "BenchmarkFileLogger"])
    rs = flags_core.define_resnet_flags()
    with pytest.raises(ValueError):
        rs.parse(["--dtype", "fp16", "--resnet_version", "1"])
    assert rs.parse([]).resnet_size == "50" and rs.parse([]).resnet_version == "2"
    assert flags_core.get_num_gpus(fs.parse(["--num_gpus", "4"])) == 4


# ------------------------------------------------------------------ loggers (logger_test.py)
def test_benchmark_loggers(tmp_path):
    assert isinstance(logger.config_benchmark_logger(None), logger.BaseBenchmarkLogger)
    ns = _fs().parse(["--benchmark_logger_type", "BenchmarkFileLogger", "--benchmark_log_dir", str(tmp_path / "log")])
    with logger.benchmark_context(ns):
        log = logger.get_benchmark_logger()
        assert isinstance(log, logger.BenchmarkFileLogger)
        log.log_metric("accuracy", 0.999, global_step=1e4, extras={"name": "value"})
        log.log_metric("loss", 0.02, unit="none")
        log.log_metric("bad", "not a number")
        log.log_evaluation_result({"loss": 0.46, "accuracy": 0.95, "global_step": 207})
        log.log_run_info("resnet", "CIFAR-10-synthetic", {"batch_size": 32, "dtype": "bf16", "synthetic_data": True, "lr": 0.1})
    lines = [json.loads(l) for l in open(tmp_path / "log" / logger.METRIC_LOG_FILE_NAME)]
    assert [l["name"] for l in lines] == ["accuracy", "loss", "accuracy", "loss"]
    assert lines[0]["value"] == 0.999 and lines[0]["global_step"] == 1e4 and lines[0]["extras"] == [{"name": "name", "value": "value"}]
    assert lines[2]["global_step"] == 207
    run = json.load(open(tmp_path / "log" / logger.BENCHMARK_RUN_LOG_FILE_NAME))
    assert run["model_name"] == "resnet" and run["dataset"] == {"name": "CIFAR-10-synthetic"}
    params = {p["name"]: p for p in run["run_parameters"]}
    assert params["batch_size"]["long_value"] == 32 and params["dtype"]["string_value"] == "bf16"
    assert params["synthetic_data"]["bool_value"] == "True" and params["lr"]["float_value"] == 0.1
    assert run["machine_config"]["cpu_info"]["num_cores"] >= 1 and "framework_version" in run
    with pytest.raises(ValueError):
        logger.config_benchmark_logger(type("F", (), {"benchmark_logger_type": "nope"})())
    with pytest.raises(RuntimeError):
        with logger.benchmark_context(None):
            raise RuntimeError("failure status path")
    assert cloud_lib.on_gcp() is False          # no network probe unless explicitly enabled


# ------------------------------------------------------------------ hooks (hooks_test.py, metric_hook_test.py)
def test_examples_per_second_hook_every_n_steps():
    with pytest.raises(ValueError):
        hooks.ExamplesPerSecondHook(256, every_n_steps=10, every_n_secs=20)
    with pytest.raises(ValueError):
        hooks.ExamplesPerSecondHook(256)
    for warm in (0, 10):
        m = mock_lib.MockBenchmarkLogger()
        h = hooks.ExamplesPerSecondHook(256, every_n_steps=10, warm_steps=warm, metric_logger=m)
        for step in range(1, 41):
            h.after_step(step, {})
        steps = sorted({r["global_step"] for r in m.logged_metric})
        assert steps == ([11, 21, 31] if warm == 0 else [20, 30, 40])
        assert {r["name"] for r in m.logged_metric} == {"average_examples_per_sec", "current_examples_per_sec"}
        assert all(r["value"] > 0 for r in m.logged_metric)


def test_examples_per_second_hook_every_n_secs():
    m = mock_lib.MockBenchmarkLogger()
    h = hooks.ExamplesPerSecondHook(64, every_n_secs=0.05, metric_logger=m)
    h.after_step(1, {})
    h.after_step(2, {})
    assert m.logged_metric == []
    time.sleep(0.06)
    h.after_step(3, {})
    assert len(m.logged_metric) == 2 and m.logged_metric[0]["global_step"] == 3


def test_logging_hooks_and_factory(tmp_path):
    with pytest.raises(ValueError):
        hooks.LoggingMetricHook(["loss"])
    m = mock_lib.MockBenchmarkLogger()
    h = hooks.LoggingMetricHook(["cross_entropy", "missing"], metric_logger=m, every_n_iter=5, at_end=True)
    for step in range(1, 13):
        h.after_step(step, {"cross_entropy": torch.tensor(1.0 / step), "learning_rate": 0.1})
    h.end()
    assert [r["global_step"] for r in m.logged_metric] == [1, 6, 11, 12]
    assert m.logged_metric[1]["value"] == pytest.approx(1 / 6)
    lines = []
    t = hooks.LoggingTensorHook(["learning_rate", "cross_entropy"], every_n_iter=2, sink=lines.append)
    for step in range(1, 6):
        t.after_step(step, {"learning_rate": 0.1, "cross_entropy": 2.0})
    assert len(lines) == 3 and "learning_rate = 0.1" in lines[0]
    p = hooks.ProfilerHook(str(tmp_path), save_steps=4)
    for step in range(1, 9):
        p.after_step(step, {})
    trace = json.load(open(tmp_path / "timeline-4.json"))
    assert len(trace["traceEvents"]) == 3 and trace["traceEvents"][0]["name"] == "train_step"
    assert hooks.get_train_hooks([]) == []
    hs = hooks.get_train_hooks(["LoggingTensorHook", "profilerhook", "ExamplesPerSecondHook", "LoggingMetricHook"],
                               model_dir=str(tmp_path), batch_size=64)
    assert [type(x).__name__ for x in hs] == ["LoggingTensorHook", "ProfilerHook", "ExamplesPerSecondHook", "LoggingMetricHook"]
    with pytest.raises(ValueError):
        hooks.get_train_hooks(["StepCounterHook"])


# ------------------------------------------------------------------ misc (distribution_utils_test.py, model_helpers_test.py)
def test_distribution_strategy_and_batch_split():
    assert distribution_utils.get_distribution_strategy(0).kind == "cpu"
    one = distribution_utils.get_distribution_strategy(1)
    assert one.kind == "one_device" and one.device == "cuda:0"
    dp = distribution_utils.get_distribution_strategy(4)
    assert (dp.kind, dp.num_gpus, dp.all_reduce_alg, dp.num_packs) == ("data_parallel", 4, "fused_p2p", 4)
    assert distribution_utils.get_distribution_strategy(2, "nccl").all_reduce_alg == "nccl"
    with pytest.raises(ValueError):
        distribution_utils.get_distribution_strategy(3)
    assert distribution_utils.per_device_batch_size(147, 1) == 147
    assert distribution_utils.per_device_batch_size(128, 4) == 32
    with pytest.raises(ValueError, match="try --batch_size=144"):
        distribution_utils.per_device_batch_size(147, 8)


def test_model_helpers(tmp_path):
    assert model_helpers.past_stop_threshold(None, -1.5) is False
    assert model_helpers.past_stop_threshold(0.54, 0.54) and model_helpers.past_stop_threshold(0.54, 0.9)
    assert not model_helpers.past_stop_threshold(0.54, 0.53)
    with pytest.raises(ValueError):
        model_helpers.past_stop_threshold("str", 1)
    with pytest.raises(ValueError):
        model_helpers.past_stop_threshold(0.5, None)
    it = model_helpers.generate_synthetic_data([5], 123, torch.float32, [], 456, torch.int32)
    for _ in range(3):
        x, y = next(it)
        assert x.tolist() == [123.0] * 5 and int(y) == 456
    x, y = next(model_helpers.generate_synthetic_data([32, 32, 3], batch_size=4))
    assert x.shape == (4, 32, 32, 3) and float(x.abs().sum()) == 0 and y.tolist() == [0, 0, 0, 0]
    d = tmp_path / "model"
    d.mkdir()
    model_helpers.apply_clean(type("F", (), {"clean": False, "model_dir": str(d)})())
    assert d.exists()
    model_helpers.apply_clean(type("F", (), {"clean": True, "model_dir": str(d)})())
    assert not d.exists()


# ------------------------------------------------------------------ export (export_test.py)
def test_export_receiver_and_bundle(tmp_path):
    from distributedtf_b200.utils.export import export
    spec = export.build_tensor_serving_input_receiver_fn([4, 5], batch_size=1)()
    assert spec["shape"] == [1, 4, 5] and spec["dtype"] == "float32" and spec["placeholder"].shape == (1, 4, 5)
    spec = export.build_tensor_serving_input_receiver_fn([4, 5], dtype=torch.int32, batch_size=None)()
    assert spec["shape"] == [None, 4, 5] and spec["dtype"] == "int32"
    from distributedtf_b200.models import resnet
    cfg = resnet.cifar10_config(8)
    params = resnet.init_params(cfg, "he_init", 0)
    import dataclasses
    export.export_inference_bundle(str(tmp_path / "exp"), params, "resnet", dataclasses.asdict(cfg), [32, 32, 3])
    predict = export.load_inference_bundle(str(tmp_path / "exp"))
    x = torch.randn(3, 32, 32, 3)
    out = predict(x)
    want = resnet.forward(cfg, params, x.permute(0, 3, 1, 2), False)
    assert torch.equal(out["classes"], want.argmax(1))
    torch.testing.assert_close(out["probabilities"], torch.softmax(want, 1))


# ------------------------------------------------------------------ testing helpers (reference_data_test.py)
def test_reference_data_harness(tmp_path):
    rd = reference_data.ReferenceData(str(tmp_path), "dense")
    gen = torch.Generator().manual_seed(reference_data.name_to_seed("dense"))
    inputs = {"x": torch.rand(5, 3, generator=gen), "w": torch.rand(3, 4, generator=gen)}
    saved = rd.save(inputs, lambda x, w: [x @ w])
    assert saved[:2] == [5, 3] and len(saved) == (2 + 3) * 3
    rd.check(lambda x, w: [x @ w])
    with pytest.raises(AssertionError):
        rd.check(lambda x, w: [x @ w + 1e-3])                     # designed numerical failure
    with pytest.raises(AssertionError):
        rd.check(lambda x, w: [x @ w, x])                         # designed structural failure
    assert reference_data.common_tensor_properties(np.arange(6).reshape(2, 3)) == [2, 3, 0.0, 5.0, 15.0]
    assert reference_data.name_to_seed("a") != reference_data.name_to_seed("b")


def test_run_synthetic_cleans_up(tmp_path):
    seen = {}

    def main(argv):
        seen["argv"] = list(argv)
        seen["dir"] = argv[argv.index("--model_dir") + 1]
        assert os.path.isdir(seen["dir"])
        return 7
    assert integration.run_synthetic(main, str(tmp_path), extra_flags=["-rs", "8"], max_train=1) == 7
    assert "--use_synthetic_data" in seen["argv"] and seen["argv"][-2:] == ["--max_train_steps", "1"]
    assert not os.path.exists(seen["dir"])
    with pytest.raises(RuntimeError):
        integration.run_synthetic(lambda a: (_ for _ in ()).throw(RuntimeError("x")), str(tmp_path))
    assert os.listdir(tmp_path) == []


# ------------------------------------------------------------------ dead-in-reference utilities
def test_file_io_round_trip(tmp_path):
    import pandas as pd
    from distributedtf_b200.utils.data import file_io
    n = 2500
    df = pd.DataFrame({"a": np.arange(n, dtype=np.int64), "b": np.linspace(0, 1, n).astype(np.float32),
                       "c": (np.arange(n) % 7).astype(np.int32)})
    path = file_io.write_to_buffer(df, str(tmp_path / "buf.records"), ["a", "b"], rows_per_core=300)
    back = file_io.read_buffer(path)
    assert set(back) == {"a", "b"}
    np.testing.assert_array_equal(back["a"], df["a"].to_numpy())
    np.testing.assert_array_equal(back["b"], df["b"].to_numpy())
    size = os.stat(path).st_size
    assert file_io.write_to_buffer(df, path, ["a", "b"], expected_size=size) == path      # reused
    tmp = file_io.write_to_temp_buffer(df, str(tmp_path / "t"), ["c"])
    assert os.path.exists(tmp)
    file_io._GC.purge()
    assert not os.path.exists(tmp)
    assert sum(len(s) for shards in file_io.iter_shard_dataframe(df, 100) for s in shards) == n


def test_embedding_matmul():
    from distributedtf_b200.utils.accelerator import tpu
    table = torch.arange(12, dtype=torch.float32).reshape(4, 3)
    values = torch.tensor([[0, 3, 1], [2, 2, 0]])
    mask = torch.tensor([[1., 1., 0.], [1., 0., 1.]])
    out = tpu.embedding_matmul(table, values, mask)
    assert torch.equal(out, table[values] * mask.unsqueeze(-1))
    with pytest.raises(ValueError):
        tpu.embedding_matmul(table, values[0], mask[0])
    with pytest.raises(NotImplementedError):
        tpu.construct_scalar_host_call({}, "/tmp")


def test_bigquery_logger_uploads_from_threads(tmp_path):
    """C12: BenchmarkBigQueryLogger — run info, RUNNING status, metrics, final status, through the uploader API;
    offline the uploader spools JSON lines per table."""
    import json
    import types
    from distributedtf_b200.utils.logs import logger
    flags = types.SimpleNamespace(benchmark_logger_type="BenchmarkBigQueryLogger", benchmark_log_dir=str(tmp_path),
                                  gcp_project=None, bigquery_data_set="ds", bigquery_run_table="run",
                                  bigquery_run_status_table="status", bigquery_metric_table="metric")
    with logger.benchmark_context(flags):
        lg = logger.get_benchmark_logger()
        assert isinstance(lg, logger.BenchmarkBigQueryLogger)
        lg.log_run_info("resnet", "cifar10", {"batch_size": 128}, test_id="t1")
        lg.log_metric("accuracy", 0.5, global_step=10)
        lg.log_metric("bad", "not a number")                 # dropped, as in the reference
        lg.log_evaluation_result({"loss": 1.25, "global_step": 20})
    spool = tmp_path / "bigquery_spool"
    rows = {f: [json.loads(l) for l in open(spool / f)] for f in sorted(os.listdir(spool))}
    assert sorted(rows) == ["ds.metric.jsonl", "ds.run.jsonl", "ds.status.jsonl"]
    assert [r["status"] for r in rows["ds.status.jsonl"]] == ["running", "success"]
    assert {r["name"] for r in rows["ds.metric.jsonl"]} == {"accuracy", "loss"}
    assert len({r["run_id"] for r in rows["ds.metric.jsonl"]}) == 1
    assert rows["ds.run.jsonl"][0]["model_name"] == "resnet" and rows["ds.run.jsonl"][0]["test_id"] == "t1"
    logger.config_benchmark_logger()                          # back to the default for later tests


# --------------------------------------------------------------------------- harness wiring
def test_cli_is_built_from_the_flag_groups():
    """The PBT command line takes the harness flags of the reference's groups (either spelling) and runs their
    validators; PBT-level arguments sit next to them."""
    from distributedtf_b200.config import build_parser, parse_args
    cfg = parse_args(["6", "--model", "cifar10", "--stop_threshold", "0.5", "--export-dir", "/tmp/e",
                      "--epochs_between_evals", "2", "--dtype", "fp16", "--loss_scale", "64",
                      "--benchmark_logger_type", "BenchmarkFileLogger", "--benchmark_log_dir", "/tmp/b"])
    assert (cfg.population_size, cfg.stop_threshold, cfg.export_dir, cfg.epochs_between_evals) == (6, 0.5, "/tmp/e", 2)
    assert cfg.dtype == "fp16" and cfg.loss_scale == 64 and cfg.benchmark_log_dir == "/tmp/b"
    assert "stop_threshold" in build_parser()._dtf_flagset.key_flags
    with pytest.raises(ValueError, match="loss_scale"):
        parse_args(["--loss_scale", "-3"])
    with pytest.raises(ValueError, match="benchmark_log_dir"):
        parse_args(["--benchmark_logger_type", "BenchmarkFileLogger"])
    assert parse_args(["--exploit-path", "p2p", "--model", "cifar10"]).scoreboard is True
    assert parse_args(["--exploit-path", "p2p", "--model", "cifar10", "--no-scoreboard"]).scoreboard is False
    assert parse_args([]).scoreboard is False


def test_run_synthetic_drives_the_cifar_harness(tmp_path):
    """integration.run_synthetic -> resnet_run.main(argv): one tiny member through the stand-alone run loop
    (flag groups, apply_clean, strategy selection, train -> eval cycle, CSV row, export)."""
    from distributedtf_b200.models import resnet_run
    from distributedtf_b200.utils.testing import integration
    export = tmp_path / "export"
    seen = {}

    def main(argv):
        seen["argv"] = list(argv)
        model_dir = argv[argv.index("--model_dir") + 1]
        out = resnet_run.main(argv)
        seen["csv"] = open(os.path.join(model_dir, "model_0", "learning_curve.csv")).read()
        return out

    acc, mid = integration.run_synthetic(main, tmp_root=str(tmp_path), max_train=2, extra_flags=[
        "--resnet_size", "8", "--batch_size", "16", "--train_epochs", "2", "--epochs_between_evals", "1",
        "--engine", "eager", "--dtype", "fp32", "--export_dir", str(export), "--stop_threshold", "0.0"])
    assert mid == 0 and 0.0 <= acc <= 1.0
    assert "--use_synthetic_data" in seen["argv"] and "--max_train_steps" in seen["argv"]
    # stop_threshold 0.0 is passed by the first evaluation: one cycle, one CSV row (plus the header)
    assert len(seen["csv"].strip().splitlines()) == 2
    assert (export / "model_0" / "signature.json").exists()
    assert not any(p.name.startswith("tmp") and p.is_dir() for p in tmp_path.iterdir() if p.name != "export")


def test_member_train_cycles_follow_epochs_between_evals(tmp_path):
    from distributedtf_b200.hpspace import default_cifar_hparam
    from distributedtf_b200.models.cifar10 import Cifar10Model
    hp = default_cifar_hparam()
    hp["batch_size"] = 16
    m = Cifar10Model(3, hp, str(tmp_path / "model_"), engine="eager", dtype="fp32", max_train_steps=1, eval_batches=1,
                     resnet_size=8, epochs_between_evals=2, device=torch.device("cpu"))
    evals = []
    real = m.evaluate
    m.evaluate = lambda **kw: (evals.append(kw.get('final')), real(**kw))[1]
    m.train(5, 5)
    assert evals == [False, True] and m.epoches_trained == 5          # cycles of 2 and 3 epochs
    rows = open(tmp_path / "model_3" / "learning_curve.csv").read().strip().splitlines()
    assert [r.split(",")[0] for r in rows[1:]] == ["1", "4"]


def test_train_accuracy_reaches_the_logging_hook(tmp_path):
    from distributedtf_b200.hpspace import default_cifar_hparam
    from distributedtf_b200.models.cifar10 import Cifar10Model
    from distributedtf_b200.utils.logs import hooks as hooks_lib
    lines = []
    hook = hooks_lib.LoggingTensorHook(["learning_rate", "cross_entropy", "train_accuracy"], every_n_iter=1,
                                       sink=lines.append)
    hp = default_cifar_hparam()
    hp["batch_size"] = 16
    m = Cifar10Model(0, hp, str(tmp_path / "model_"), engine="eager", dtype="fp32", max_train_steps=2, eval_batches=1,
                     resnet_size=8, device=torch.device("cpu"), hooks=[hook])
    m.train(1, 1)
    assert len(lines) == 2 and all("train_accuracy = " in ln and "cross_entropy = " in ln for ln in lines)
