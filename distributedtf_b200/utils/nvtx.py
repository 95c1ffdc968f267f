"""NVTX ranges around the PBT phases, visible in Nsight Systems / Compute timelines.

``with nvtx.range("train member 3"):`` pushes a range on the calling thread when CUDA is initialised and is a
no-op otherwise (CPU runs, the toy model).  Disable with ``DTF_NVTX=0``.  The reference's only tracing aid is the
TensorFlow ``ProfilerHook`` (``official/utils/logs/hooks_helper.py:97-109``).
"""
from __future__ import annotations

import contextlib
import os

_ENABLED = os.environ.get("DTF_NVTX", "1") != "0"


def _live() -> bool:
    if not _ENABLED:
        return False
    try:
        import torch
        return torch.cuda.is_available() and torch.cuda.is_initialized()
    except Exception:
        return False


@contextlib.contextmanager
def range(name: str):       # noqa: A001  (mirrors torch.cuda.nvtx.range)
    if not _live():
        yield
        return
    import torch
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


def mark(name: str) -> None:
    if _live():
        import torch
        torch.cuda.nvtx.mark(name)
