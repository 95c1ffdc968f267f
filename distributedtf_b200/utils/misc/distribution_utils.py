"""Distribution-strategy selection.

Parity: reference ``official/utils/misc/distribution_utils.py:24-78``: 0 GPUs -> one CPU device,
1 GPU -> one device, >1 -> mirrored data parallelism with an optional all-reduce algorithm.  The
reference hard-wires this to 1 GPU (``resnet_run_loop.py:390-392``); here >1 is live and maps onto
``parallel/dp.py`` (fused one-shot all-reduce + update, or NCCL as the library baseline).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

from ...parallel.dp import per_device_batch_size  # noqa: F401  (re-export, same contract as the reference)

ALL_REDUCE_ALGS = (None, "fused_p2p", "nccl", "hierarchical_copy")


@dataclass(frozen=True)
class DistributionStrategy:
    kind: str                    # "cpu" | "one_device" | "data_parallel"
    num_gpus: int
    all_reduce_alg: Optional[str] = None
    num_packs: int = 1

    @property
    def device(self) -> str:
        return "cpu" if self.kind == "cpu" else "cuda:0"


def get_distribution_strategy(num_gpus: int, all_reduce_alg: Optional[str] = None) -> DistributionStrategy:
    if all_reduce_alg not in ALL_REDUCE_ALGS:
        raise ValueError(f"all_reduce_alg must be one of {ALL_REDUCE_ALGS}")
    if num_gpus == 0:
        return DistributionStrategy("cpu", 0)
    if num_gpus == 1:
        return DistributionStrategy("one_device", 1)
    if num_gpus not in (2, 4, 8):
        raise ValueError("a data-parallel member spans 2, 4 or 8 GPUs")
    return DistributionStrategy("data_parallel", num_gpus, all_reduce_alg or "fused_p2p", num_packs=num_gpus)
