"""Run reports: per-member CSV curves, JSON dumps, optional PNG plots.

Parity: reference ``pbt_cluster.py:240-470`` and ``main_manager.py:60-61``.  The
CSV *column order is API* (the plots index columns by position: col0 step/epoch,
col1 accuracy, col3 learning rate — ``pbt_cluster.py:329,370,413``).  matplotlib
is optional here: curve data and the top-3 average are always written as
CSV/JSON; PNGs are produced with the Agg CPU rasteriser only when matplotlib
imports (no graphics stack is needed or probed).
"""
from __future__ import annotations

import csv
import json
import os
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple


def mode_suffix(do_exploit: bool, do_explore: bool) -> Tuple[str, str]:
    if do_exploit and do_explore:
        return "PBT", "PBT"
    if do_exploit:
        return "Exploit only", "exploit_only"
    if do_explore:
        return "Explore only", "explore_only"
    return "Grid search", "grid_search"


def append_csv(path: str, fields: Sequence[str], rows: Iterable[Dict[str, Any]]) -> None:
    """Append rows; the header is written only when the file is created."""
    new = not os.path.isfile(path)
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "a", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(fields))
        if new:
            w.writeheader()
        for r in rows:
            w.writerow(r)


def dump_population(path: str, all_values: List[List[Any]]) -> None:
    """``initial_hp.json``: list of {model_id, accuracy, hparams} sorted by accuracy."""
    ordered = sorted(all_values, key=lambda v: v[1])
    doc = [{"model_id": v[0], "accuracy": float(v[1]), "hparams": v[2]} for v in ordered]
    with open(path, "w") as f:
        json.dump(doc, f, indent=4, sort_keys=True)


def dump_best(path: str, all_values: List[List[Any]]) -> Dict[str, Any]:
    if not all_values:                     # every member was dropped (NaN / exception)
        doc = {"best_model_id": None, "best_acc": float("nan"), "best_hparams": None}
        with open(path, "w") as f:
            json.dump(doc, f, indent=4, sort_keys=True)
        return doc
    best = sorted(all_values, key=lambda v: v[1])[-1]
    doc = {"best_model_id": best[0], "best_acc": float(best[1]), "best_hparams": best[2]}
    with open(path, "w") as f:
        json.dump(doc, f, indent=4, sort_keys=True)
    return doc


def append_timing(path: str, n: int, pop_size: int, seconds: float) -> None:
    with open(path, "a") as f:
        f.write("n = {}, pop_size = {}, time = {}s\n".format(n, pop_size, seconds))


# --------------------------------------------------------------------------- curves
def _member_dirs(save_root: str) -> List[str]:
    if not os.path.isdir(save_root):
        return []
    return sorted(os.path.join(save_root, d) for d in os.listdir(save_root) if d.startswith("model_"))


def read_curves(save_root: str, column: int, filename: str = "learning_curve.csv"
                ) -> List[List[Tuple[int, float]]]:
    """[(col0 as int, col<column> as float)] per member, by column *position*."""
    curves = []
    for d in _member_dirs(save_root):
        p = os.path.join(d, filename)
        if not os.path.isfile(p):
            continue
        with open(p, newline="") as f:
            rd = csv.reader(f)
            header = next(rd, None)
            if header is None:
                continue
            pts = [(int(float(row[0])), float(row[column])) for row in rd if row]
        curves.append(pts)
    return curves


def read_theta(save_root: str) -> List[List[Tuple[float, float]]]:
    out = []
    for d in _member_dirs(save_root):
        p = os.path.join(d, "theta.csv")
        if not os.path.isfile(p):
            continue
        with open(p, newline="") as f:
            rd = csv.reader(f)
            next(rd, None)
            out.append([(float(r[0]), float(r[1])) for r in rd if r])
    return out


def top3_average(curves: List[List[Tuple[int, float]]]) -> List[Tuple[int, float]]:
    """Per record index: mean of the three best accuracies (fewer if < 3 curves)."""
    longest = max((len(c) for c in curves), default=0)
    out: List[Tuple[int, float]] = []
    for i in range(longest):
        column = sorted(c[i][1] for c in curves if len(c) > i)
        epoch = next((c[i][0] for c in curves if len(c) > i), 0)
        if not column:
            out.append((epoch, 0.0))
        elif len(column) < 3:
            out.append((epoch, sum(column) / len(column)))
        else:
            out.append((epoch, sum(column[-3:]) / 3.0))
    return out


def _pyplot():
    try:
        import matplotlib
        matplotlib.use("Agg")
        from matplotlib import pyplot
        return pyplot
    except Exception:
        return None


def _finish(pyplot, save_root: str, stem: str, title: str, suffix: str) -> str:
    out = os.path.join(save_root, f"{stem}_{suffix}.png")
    pyplot.title(title)
    pyplot.savefig(out)
    pyplot.close()
    return out


def write_reports(save_root: str, do_exploit: bool, do_explore: bool, toy: bool = False
                  ) -> Dict[str, Optional[str]]:
    """Emit every report the reference produces; returns {name: path-or-None}."""
    title, suffix = mode_suffix(do_exploit, do_explore)
    produced: Dict[str, Optional[str]] = {}
    acc = read_curves(save_root, 1)
    lr = read_curves(save_root, 3)
    avg = top3_average(acc)
    summary = os.path.join(save_root, f"curves_{suffix}.json")
    with open(summary, "w") as f:
        json.dump({"mode": title, "accuracy": acc, "lr": lr, "best3": avg}, f)
    produced["curves_json"] = summary
    plt = _pyplot()
    for name in ("toy", "acc", "lr", "best3"):
        produced[name] = None
    if plt is None:
        return produced
    if toy:
        thetas = read_theta(save_root)
        import numpy as np
        xs, ys = np.meshgrid(np.linspace(0, 1, 100), np.linspace(0, 1, 100))
        plt.figure()
        plt.xlabel(r"$\theta_0$"); plt.ylabel(r"$\theta_1$"); plt.xlim(0, 1); plt.ylim(0, 1)
        for t in thetas:
            if t:
                plt.plot([p[0] for p in t], [p[1] for p in t], ".")
        plt.contour(xs, ys, 1.2 - (xs ** 2 + ys ** 2), colors="lightgray")
        produced["toy"] = _finish(plt, save_root, "toy", title, suffix)
    plt.figure()
    for c in acc:
        plt.plot([p[0] for p in c], [p[1] for p in c])
    plt.xlabel("Train epochs"); plt.ylabel("Accuracy"); plt.grid(True)
    produced["acc"] = _finish(plt, save_root, "acc", title, suffix)
    plt.figure()
    for c in lr:
        plt.plot([p[0] for p in c], [p[1] for p in c])
    plt.xlabel("Train epochs"); plt.ylabel("Learning rate"); plt.ylim(0, 1); plt.grid(True)
    produced["lr"] = _finish(plt, save_root, "lr", title, suffix)
    plt.figure()
    for c in acc:
        plt.plot([p[0] for p in c], [p[1] for p in c], color=(0.0, 0.0, 0.5, 0.3))
    if avg:
        plt.plot([p[0] for p in avg], [p[1] for p in avg], "r")
    plt.xlabel("Train epochs"); plt.ylabel("Accuracy"); plt.ylim(0, 1); plt.grid(True)
    produced["best3"] = _finish(plt, save_root, "best3", title, suffix)
    return produced
