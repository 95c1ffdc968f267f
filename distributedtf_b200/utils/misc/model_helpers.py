"""Misc model helpers (reference ``official/utils/misc/model_helpers.py:27-93``)."""
from __future__ import annotations

import numbers
import os
import shutil
from typing import Iterator, Optional, Sequence, Tuple

import torch


def past_stop_threshold(stop_threshold: Optional[float], eval_metric) -> bool:
    """True when ``eval_metric >= stop_threshold``; ``None`` threshold never stops."""
    if stop_threshold is None:
        return False
    if not isinstance(stop_threshold, numbers.Number):
        raise ValueError("Threshold for checking stop conditions must be a number.")
    if not isinstance(eval_metric, numbers.Number):
        raise ValueError("Eval metric being checked against stop conditions must be a number.")
    return eval_metric >= stop_threshold


def generate_synthetic_data(input_shape: Sequence[int], input_value=0, input_dtype=torch.float32,
                            label_shape: Sequence[int] = (), label_value=0, label_dtype=torch.int32,
                            batch_size: Optional[int] = None, device="cpu"
                            ) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
    """Endless stream of one constant (features, labels) element, optionally batched — the reference's
    synthetic dataset (zeros images, label 0 by default)."""
    shape_x = ([batch_size] if batch_size else []) + list(input_shape)
    shape_y = ([batch_size] if batch_size else []) + list(label_shape)
    x = torch.full(shape_x, input_value, dtype=input_dtype, device=device)
    y = torch.full(shape_y, label_value, dtype=label_dtype, device=device)
    while True:
        yield x, y


def apply_clean(flags_obj) -> None:
    if getattr(flags_obj, "clean", False) and os.path.exists(flags_obj.model_dir):
        shutil.rmtree(flags_obj.model_dir)
