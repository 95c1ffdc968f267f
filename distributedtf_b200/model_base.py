"""The contract every population member implements.

Parity: reference ``model_base.py:11-113``.  Same attribute and method names so a
model written against the reference drops in: ``cluster_id, hparams,
save_base_dir, epoches_trained, need_explore, accuracy``; ``train(num_epoch,
total_epochs)`` (abstract), ``perturb_hparams()``, ``get_accuracy()``,
``get_values() -> [id, accuracy, hparams]``, ``set_values(values)``.

Differences by design (B200-first):

* members are *persistent*: they are built once and keep their state on the GPU
  between rounds; the checkpoint on disk (``save()`` / ``restore()``) is written at
  round end for resume and for the file-copy exploit baseline, not re-read every
  round;
* ``set_values`` may receive an extended record ``[id, acc, hparams, extras]``
  where ``extras`` names the winner (``{"src_id", "src_rank"}``) so the weight
  transfer can run device-to-device (``parallel/exploit.py``);
* ``drop()`` removes the member's directory (worker member-drop semantics,
  reference ``training_worker.py:63-80``).
"""
from __future__ import annotations

import copy
import os
import random
import shutil
from typing import Any, Dict, List, Optional, Sequence

from . import explore as _explore


class ModelBase:
    def __init__(self, cluster_id: int, hparams: Dict[str, Any], save_base_dir: str):
        self.cluster_id = cluster_id
        self.hparams = hparams
        self.save_base_dir = save_base_dir
        self.epoches_trained = 0
        self.need_explore = False
        self._perturb_factors = [0.8, 1.2]
        self._rng: Optional[random.Random] = None
        bs = self.hparams.get("batch_size")
        if bs is not None and not isinstance(bs, int):
            # numpy scalars / 0-d arrays from samplers
            self.hparams["batch_size"] = int(getattr(bs, "item", lambda: bs)())
        self.accuracy = 0.0

    # ------------------------------------------------------------------ contract
    def train(self, num_epoch: int, total_epochs: int) -> None:
        """Train ``num_epoch`` epochs, evaluate, checkpoint, append the CSV row(s)."""
        raise NotImplementedError

    def perturb_hparams(self) -> None:
        _explore.perturb_hparams(self.hparams, self._rng, tuple(self._perturb_factors))
        self.on_hparams_changed()

    def get_accuracy(self) -> float:
        return self.accuracy

    def get_values(self) -> List[Any]:
        return [self.cluster_id, self.get_accuracy(), self.hparams]

    def set_values(self, values: Sequence[Any]) -> None:
        self.hparams = copy.deepcopy(values[2])
        extras = values[3] if len(values) > 3 else None
        if extras:
            self.adopt_state(extras)
        self.on_hparams_changed()

    # ------------------------------------------------------------------ extensions
    @property
    def save_dir(self) -> str:
        return self.save_base_dir + str(self.cluster_id)

    def seed(self, seed: int) -> None:
        self._rng = random.Random(seed)

    def on_hparams_changed(self) -> None:
        """Hook: push the new hyper-parameters to the device-resident block."""

    def adopt_state(self, extras: Dict[str, Any]) -> None:
        """Hook: exploit — take over the winner's weights / optimizer state."""

    def save(self) -> None:
        """Hook: write the member checkpoint into ``save_dir``."""

    def restore(self) -> bool:
        """Hook: load the member checkpoint; returns False if none exists."""
        return False

    def drop(self) -> None:
        shutil.rmtree(self.save_dir, ignore_errors=True)

    def ensure_dir(self) -> str:
        os.makedirs(self.save_dir, exist_ok=True)
        return self.save_dir
