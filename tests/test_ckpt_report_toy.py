"""Checkpoint layout, file-copy exploit, reports, toy model."""
import csv
import json
import os

import numpy as np
import pytest

from distributedtf_b200 import ckpt, report
from distributedtf_b200.config import RunConfig, parse_args
from distributedtf_b200.launch import run_local
from distributedtf_b200.models import toy


def test_checkpoint_round_trip(tmp_path):
    d = str(tmp_path / "model_0")
    import torch
    tensors = {"w": np.arange(10, dtype=np.float32).reshape(2, 5),
               "slot": torch.arange(7, dtype=torch.float32),
               "wb": torch.tensor([1.5, -2.0, 3.25], dtype=torch.bfloat16),
               "step": np.asarray([42], dtype=np.int64)}
    ckpt.save_checkpoint(d, tensors, step=42, optimizer="Momentum", meta={"k": 1})
    assert sorted(os.listdir(d)) == ["checkpoint", "model.ckpt.data", "model.ckpt.index"]
    assert 'model_checkpoint_path: "model.ckpt"' in open(os.path.join(d, "checkpoint")).read()
    out, index = ckpt.load_checkpoint(d)
    assert index["step"] == 42 and index["optimizer"] == "Momentum" and index["meta"] == {"k": 1}
    np.testing.assert_array_equal(out["w"], tensors["w"])
    np.testing.assert_array_equal(out["slot"], np.arange(7, dtype=np.float32))
    assert torch.from_numpy(out["wb"]).view(torch.bfloat16).tolist() == [1.5, -2.0, 3.25]
    assert all(e["offset"] % ckpt.ALIGN == 0 for e in index["entries"].values())
    assert ckpt.load_checkpoint(str(tmp_path / "nope")) is None


def test_file_copy_exploit_keeps_history_files(tmp_path):
    src, dst = tmp_path / "model_1", tmp_path / "model_0"
    for d in (src, dst):
        d.mkdir()
        (d / "eval").mkdir()
    ckpt.save_checkpoint(str(src), {"w": np.ones(4, np.float32)}, step=5)
    ckpt.save_checkpoint(str(dst), {"w": np.zeros(4, np.float32)}, step=2)
    for name, text in [("learning_curve.csv", "mine"), ("theta.csv", "mine"),
                       ("events.out.tfevents.1", "mine"), (".nfs123", "mine"), ("stale.bin", "old")]:
        (dst / name).write_text(text)
    for name in ("learning_curve.csv", "theta.csv", "events.out.tfevents.9", "graph.pbtxt"):
        (src / name).write_text("winner")
    (src / "eval" / "x").write_text("sub")
    ckpt.copy_member_files(str(src), str(dst))
    assert (dst / "learning_curve.csv").read_text() == "mine"
    assert (dst / "theta.csv").read_text() == "mine"
    assert (dst / "events.out.tfevents.1").read_text() == "mine" and not (dst / "events.out.tfevents.9").exists()
    assert not (dst / "stale.bin").exists() and (dst / "graph.pbtxt").read_text() == "winner"
    assert not (dst / "eval" / "x").exists()
    out, index = ckpt.load_checkpoint(str(dst))
    assert index["step"] == 5 and out["w"].tolist() == [1, 1, 1, 1]     # bit-identical to the winner
    assert (dst / "model.ckpt.data").read_bytes() == (src / "model.ckpt.data").read_bytes()


def test_toy_save_restore_step_counts(tmp_path):
    """Same scenario as reference test_toy_model.py:38-50."""
    base = str(tmp_path / "savedata" / "model_")
    os.makedirs(tmp_path / "savedata")
    hp = {"h_0": 1.0, "h_1": 0.0, "opt_case": {"optimizer": "gd", "lr": 0.02}}
    assert toy.main(hp, 0, base, "", 10)[0] == 10
    assert toy.main(hp, 0, base, "", 10)[0] == 20      # restored
    assert toy.main(hp, 1, base, "", 10)[0] == 10      # new id starts fresh
    import shutil
    shutil.rmtree(tmp_path / "savedata"); os.makedirs(tmp_path / "savedata")
    assert toy.main(hp, 0, base, "", 10)[0] == 10      # wiped -> fresh


def test_toy_dynamics_and_model_contract(tmp_path):
    base = str(tmp_path / "model_")
    m0 = toy.ToyModel(0, {"h_0": 0.5, "h_1": 0.5, "opt_case": {"optimizer": "gd", "lr": 0.1}, "batch_size": 100}, base)
    m1 = toy.ToyModel(1, {"h_0": 0.5, "h_1": 0.5, "opt_case": {"optimizer": "gd", "lr": 0.1}, "batch_size": 100}, base)
    assert (m0.hparams["h_0"], m0.hparams["h_1"]) == (0.0, 1.0)
    assert (m1.hparams["h_0"], m1.hparams["h_1"]) == (1.0, 0.0)
    a0 = m1.accuracy
    m1.train(1, 1)
    assert m1.accuracy != a0 and m1.epoches_trained == 1
    # member 1 only sees theta_1's curvature: theta_0 is untouched, theta_1 shrinks
    t, _ = ckpt.load_checkpoint(m1.save_dir)
    assert t["theta"][0] == pytest.approx(0.9) and t["theta"][1] < 0.9
    # hand-computed first step: d = -(0.9^2) = -0.81; grad_1 = 2*d*(2*(-1)*0.9) = 2.916; theta_1 = 0.9 - 0.02*2.916
    assert t["theta"][1] == pytest.approx(0.9 - 0.02 * 2.916, rel=1e-5)
    m1.set_values([1, 0.3, {"h_0": 0.1, "h_1": 0.9, "opt_case": {"optimizer": "Adam", "lr": 1.0}}])
    assert (m1.hparams["h_0"], m1.hparams["h_1"]) == (1.0, 0.0)       # h never copied
    assert m1.hparams["opt_case"]["optimizer"] == "gd"
    rows = list(csv.reader(open(os.path.join(m1.save_dir, "learning_curve.csv"))))
    assert rows[0] == ["global_step", "accuracy", "optimizer", "lr"]
    assert list(csv.reader(open(os.path.join(m1.save_dir, "theta.csv"))))[0] == ["theta_0", "theta_1"]


@pytest.mark.parametrize("exploit,explore,suffix", [(True, True, "PBT"), (True, False, "exploit_only"),
                                                    (False, True, "explore_only"), (False, False, "grid_search")])
def test_toy_pbt_run_and_reports(tmp_path, exploit, explore, suffix):
    cfg = RunConfig(target_model="toy", population_size=2, train_round=6, epochs_per_round=4,
                    do_exploit=exploit, do_explore=explore, workers=1, save_root=str(tmp_path / "sd"),
                    results_file=str(tmp_path / "res.txt"), verbose=False, seed=0).validate()
    out = run_local(cfg)
    sd = tmp_path / "sd"
    assert (sd / "initial_hp.json").exists() and (sd / "best_model.json").exists()
    assert (sd / f"curves_{suffix}.json").exists()
    init = json.load(open(sd / "initial_hp.json"))
    assert [set(r) for r in init] == [{"model_id", "accuracy", "hparams"}] * 2
    assert set(json.load(open(sd / "best_model.json"))) == {"best_model_id", "best_acc", "best_hparams"}
    thetas = report.read_theta(str(sd))
    assert len(thetas) == 2 and all(len(t) == 24 for t in thetas)
    if exploit:
        # with exploit theta moves on both axes (each member descends the other's axis in turn)
        assert out["best"]["best_acc"] > 0.5
        assert len(out["history"]) == 6 and all(len(h["pairs"]) == 1 for h in out["history"])
    else:
        # without exploit each member only descends one axis: objective stays below 1.2 - 0.81
        assert out["best"]["best_acc"] < 0.39
    assert (tmp_path / "res.txt").read_text().startswith("n = 2, pop_size = 2, time = ")


def test_report_helpers(tmp_path):
    for i, accs in enumerate([[0.1, 0.5], [0.2, 0.6, 0.7], [0.3, 0.1], [0.9, 0.2]]):
        report.append_csv(str(tmp_path / f"model_{i}" / "learning_curve.csv"),
                          ["epochs", "eval_accuracy", "optimizer", "learning_rate"],
                          [{"epochs": e, "eval_accuracy": a, "optimizer": "gd", "learning_rate": 0.1 * (i + 1)}
                           for e, a in enumerate(accs)])
    acc = report.read_curves(str(tmp_path), 1)
    lr = report.read_curves(str(tmp_path), 3)
    assert len(acc) == 4 and lr[1][0] == (0, pytest.approx(0.2))
    avg = report.top3_average(acc)
    assert avg[0] == (0, pytest.approx((0.9 + 0.3 + 0.2) / 3))
    assert avg[1] == (1, pytest.approx((0.6 + 0.5 + 0.2) / 3))
    assert avg[2] == (2, pytest.approx(0.7))
    assert report.mode_suffix(False, False) == ("Grid search", "grid_search")
    out = report.write_reports(str(tmp_path), True, True)
    assert os.path.isfile(out["curves_json"])


def test_cli_parsing():
    cfg = parse_args(["40", "--model", "cifar10", "--no-explore", "--workers", "8", "--exploit-path", "p2p"])
    assert cfg.population_size == 40 and cfg.target_model == "cifar10"
    assert cfg.do_explore is False and cfg.do_exploit is True and cfg.workers == 8
    with pytest.raises(ValueError):
        RunConfig(resnet_size=33).validate()
    with pytest.raises(ValueError):
        RunConfig(resnet_version=1, dtype="fp16").validate()


def test_async_checkpoint_writer_roundtrip(tmp_path):
    """Native background writer (csrc/host_io.cpp): same on-disk format as the synchronous path, ordered
    data -> index -> pointer, many jobs in flight, reusable staging slots.  Host tensors here (no GPU)."""
    import numpy as np
    import torch
    from distributedtf_b200 import ckpt
    from distributedtf_b200.ops import native
    if not native.available():
        pytest.skip("native library not built")
    w = ckpt.AsyncCheckpointWriter(threads=3, staging_slots=2)
    states = []
    for m in range(7):                                        # more jobs than staging slots
        state = torch.arange(4096 + 256 * m, dtype=torch.int32).view(torch.uint8).clone()
        f = state.view(torch.float32)
        views = {"params": f[:1000].view(10, 100), "slot0": f[1000:2000], "step": state[8000:8008].view(torch.int64)}
        t = w.save_state(str(tmp_path / f"model_{m}"), state, views, step=m + 5, optimizer="Momentum", meta={"m": m})
        assert t > 0
        states.append((state, views))
    w.wait_all()
    assert w.pending() == 0 and w.bytes_written() == sum(s.numel() for s, _ in states)
    for m, (state, views) in enumerate(states):
        d = tmp_path / f"model_{m}"
        assert sorted(os.listdir(d)) == ["checkpoint", "model.ckpt.data", "model.ckpt.index"]      # no temp files left
        tensors, index = ckpt.load_checkpoint(str(d))
        assert index["step"] == m + 5 and index["optimizer"] == "Momentum" and index["meta"] == {"m": m}
        for k, v in views.items():
            assert np.array_equal(tensors[k], v.numpy()), k
    w.close()
