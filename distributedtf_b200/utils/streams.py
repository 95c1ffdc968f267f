"""Streams that are really new.

``torch.cuda.Stream()`` returns streams from a pool of 32 per device, round-robin: the 33rd stream a process asks for
is the first one again.  A PBT worker with dozens of members (each with a compute stream, a side stream and a
graph-capture stream) runs through that pool several times over, so one member's capture stream was another
member's compute stream or the master's score-board polling stream -- kernels of the one were recorded into, or
rejected by, the CUDA-graph capture of the other (``cudaErrorStreamCaptureUnsupported``, invalidated captures, and
the flaky concurrent-member runs of round 2).  ``new_stream`` creates a CUDA stream of its own through the native
library and wraps it as ``torch.cuda.ExternalStream``.
"""
from __future__ import annotations

import ctypes as C

import torch


def new_stream(device) -> torch.cuda.Stream:
    """A non-blocking CUDA stream of its own, wrapped for torch.  It lives until ``release`` is called on it (engines
    do that in ``close()``) or the process ends: there is deliberately no finalizer, because a stream destroyed from a
    garbage-collection callback during interpreter shutdown runs into a CUDA runtime that is already being torn down
    (it crashed pytest processes after all tests had passed)."""
    import os
    from ..ops import native
    dev = torch.device(device)
    if os.environ.get("DTF_POOLED_STREAMS") == "1":        # escape hatch: torch's pooled streams
        return torch.cuda.Stream(device=dev)
    index = dev.index if dev.index is not None else torch.cuda.current_device()
    lib = native.lib()
    handle = C.c_void_p()
    lib.dtf_stream_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    native.check(lib.dtf_stream_create(index, C.byref(handle)), "dtf_stream_create")
    stream = torch.cuda.ExternalStream(handle.value, device=torch.device("cuda", index))
    stream._dtf_owned = (index, handle.value)
    return stream


def release(stream) -> None:
    """Destroy a stream made by ``new_stream`` (queued work finishes first); a no-op for anything else and during
    interpreter shutdown."""
    import sys
    owned = getattr(stream, "_dtf_owned", None)
    if owned is None or sys.is_finalizing():
        return
    from ..ops import native
    lib = native.lib()
    lib.dtf_stream_destroy.argtypes = [C.c_int, C.c_void_p]
    stream._dtf_owned = None
    lib.dtf_stream_destroy(owned[0], C.c_void_p(owned[1]))
