"""Synthetic-data integration runner (reference ``official/utils/testing/integration.py:32-65``):
run a model ``main`` for one training step on synthetic data in a throw-away directory."""
from __future__ import annotations

import os
import shutil
import tempfile
from typing import Callable, Optional, Sequence


def run_synthetic(main: Callable[[Sequence[str]], object], tmp_root: Optional[str] = None,
                  extra_flags: Optional[Sequence[str]] = None, synth: bool = True, max_train: Optional[int] = 1):
    """Calls ``main(argv)`` with ``--model_dir <tmp> [--use_synthetic_data] [--max_train_steps N]`` plus
    ``extra_flags``; the model directory is removed afterwards, even on failure."""
    model_dir = tempfile.mkdtemp(dir=tmp_root)
    args = ["--model_dir", model_dir] + list(extra_flags or [])
    if synth:
        args.append("--use_synthetic_data")
    if max_train is not None:
        args += ["--max_train_steps", str(max_train)]
    try:
        return main(args)
    finally:
        if os.path.exists(model_dir):
            shutil.rmtree(model_dir)
