"""The reference vendors TPU helpers it never calls (``official/utils/accelerator/tpu.py:27-115``).
There is no TPU path here (BASELINE: out of scope).  ``embedding_matmul`` — an embedding lookup written
as a one-hot matmul so that it maps onto matrix units — is kept because it is hardware-neutral; the
host-call summary helper has no meaning without TPUs and raises."""
from __future__ import annotations

import torch


def construct_scalar_host_call(metric_dict, model_dir, prefix=""):
    raise NotImplementedError("TPU host calls do not exist on this platform; use utils.logs hooks / "
                              "benchmark loggers for scalar metrics")


def embedding_matmul(embedding_table: torch.Tensor, values: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """``embedding_table[values] * mask`` computed as ``one_hot(values) @ embedding_table``.
    ``values``/``mask``: [batch, length]; table: [vocab, width] -> [batch, length, width]."""
    if values.dim() != 2 or mask.shape != values.shape or embedding_table.dim() != 2:
        raise ValueError("expected rank-2 values/mask of equal shape and a rank-2 table")
    one_hot = torch.nn.functional.one_hot(values.long(), embedding_table.shape[0]).to(embedding_table.dtype)
    return torch.matmul(one_hot, embedding_table) * mask.unsqueeze(-1).to(embedding_table.dtype)
