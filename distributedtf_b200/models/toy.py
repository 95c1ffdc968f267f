"""The two-parameter toy problem from the PBT paper.

Parity: reference ``toy_model.py:7-88``.  theta starts at (0.9, 0.9); the reported
"accuracy" is ``1.2 - (t0^2 + t1^2)``; each member sees only a surrogate
``1.2 - (h0 t0^2 + h1 t1^2)`` and descends ``(objective - surrogate)^2`` with plain
gradient descent, learning rate 0.02; one "epoch" is one step.  Member 0 is pinned
to h = (0, 1), every other member to h = (1, 0); exploit therefore copies
*weights only* (``set_values`` is overridden, ``toy_model.py:83-89``) while
explore may still nudge h until the next ``set_values`` — a reference quirk kept
on purpose.

This member is host-only (BASELINE config 1: "plumbing, no GPU"): two float32
scalars do not deserve a kernel launch.  State lives in the checkpoint directory
exactly like the reference, so the file-copy exploit works unchanged.
"""
from __future__ import annotations

import os
from typing import Any, Dict, Sequence, Tuple

import numpy as np

from .. import ckpt, report
from ..model_base import ModelBase

TOY_LR = np.float32(0.02)
THETA_INIT = (0.9, 0.9)


def _objective(theta: np.ndarray) -> np.float32:
    return np.float32(1.2) - (theta[0] * theta[0] + theta[1] * theta[1])


def _step(theta: np.ndarray, h: np.ndarray) -> np.ndarray:
    # d = objective - surrogate = (h0 - 1) t0^2 + (h1 - 1) t1^2 ;  loss = d^2
    coeff = h - np.float32(1.0)
    d = np.float32(np.dot(coeff, theta * theta))
    grad = np.float32(2.0) * d * (np.float32(2.0) * coeff * theta)
    return (theta - TOY_LR * grad).astype(np.float32)


def main(hp: Dict[str, Any], model_id: int, save_base_dir: str, data_dir: str,
         train_epochs: int) -> Tuple[int, float]:
    """Stateless round: restore (if a checkpoint exists) -> train -> save -> log."""
    save_dir = save_base_dir + str(model_id)
    theta = np.asarray(THETA_INIT, dtype=np.float32)
    global_step = 0
    loaded = ckpt.load_checkpoint(save_dir) if os.path.isdir(save_dir) else None
    if loaded is not None:
        tensors, index = loaded
        theta = tensors["theta"].astype(np.float32)
        global_step = int(index["step"])
    h = np.asarray([hp["h_0"], hp["h_1"]], dtype=np.float32)
    opt = hp.get("opt_case", {})
    rows = []
    for _ in range(train_epochs):
        rows.append((float(theta[0]), float(theta[1]), global_step, float(_objective(theta)),
                     opt.get("optimizer", "gd"), opt.get("lr", float(TOY_LR))))
        theta = _step(theta, h)
        global_step += 1
    ckpt.save_checkpoint(save_dir, {"theta": theta}, step=global_step, optimizer="gd",
                         meta={"model": "toy"})
    report.append_csv(os.path.join(save_dir, "theta.csv"), ["theta_0", "theta_1"],
                      ({"theta_0": r[0], "theta_1": r[1]} for r in rows))
    report.append_csv(os.path.join(save_dir, "learning_curve.csv"),
                      ["global_step", "accuracy", "optimizer", "lr"],
                      ({"global_step": r[2], "accuracy": r[3], "optimizer": r[4], "lr": r[5]}
                       for r in rows))
    return global_step, float(_objective(theta))


class ToyModel(ModelBase):
    def __init__(self, cluster_id: int, hparams: Dict[str, Any], save_base_dir: str, **_):
        super().__init__(cluster_id, hparams, save_base_dir)
        self._pin_h()
        self.global_step = 0

    def _pin_h(self) -> None:
        if self.cluster_id == 0:
            self.hparams["h_0"], self.hparams["h_1"] = 0.0, 1.0
        else:
            self.hparams["h_0"], self.hparams["h_1"] = 1.0, 0.0

    def train(self, epoches_to_train: int, total_epochs: int) -> None:
        self.global_step, self.accuracy = main(self.hparams, self.cluster_id,
                                               self.save_base_dir, "", epoches_to_train)
        self.epoches_trained += epoches_to_train

    def set_values(self, values: Sequence[Any]) -> None:
        # exploit moves theta (through the checkpoint), never h
        self._pin_h()
        extras = values[3] if len(values) > 3 else None
        if extras:
            self.adopt_state(extras)

    def adopt_state(self, extras: Dict[str, Any]) -> None:
        if extras.get("path", "file") == "file" or extras.get("src_id") == self.cluster_id:
            return  # the master already copied the files
        # Host member under a device exploit path: fall back to a worker-side file copy.
        ckpt.copy_member_files(self.save_base_dir + str(extras["src_id"]), self.save_dir)
