"""Member checkpoint layout under ``savedata/model_<id>/``.

Parity: SURVEY §5.4.  The reference leans on ``tf.train.Saver`` / Estimator
(``toy_model.py:23,29,39``; ``resnet_run_loop.py:397-398``): a ``checkpoint``
pointer file plus ``model.ckpt*`` payload holding *all* variables — weights,
optimizer slots, BN moving statistics and ``global_step`` — and the master's
exploit is a file copy of everything except the CSV logs
(``pbt_cluster.py:168-181``).

Here the payload is a single flat dump of the member's arena:

    checkpoint            text pointer:  model_checkpoint_path: "model.ckpt"
    model.ckpt.index      JSON: {"version", "step", "optimizer", "entries":
                                 {name: {"offset","shape","dtype","nbytes"}}, "meta": {...}}
    model.ckpt.data       raw little-endian bytes, entries at 256-byte aligned offsets

so the file-copy exploit (the baseline path) keeps working unchanged, and one
device->host copy of the arena is the whole save.  The native writer
(``csrc/host_io.cpp``) is used when the library is built; this module is the
portable implementation and the reader.
"""
from __future__ import annotations

import json
import os
import shutil
import tempfile
from typing import Any, Dict, Mapping, Optional, Tuple

import numpy as np

CKPT_PREFIX = "model.ckpt"
POINTER_FILE = "checkpoint"
ALIGN = 256
FORMAT_VERSION = 1

# Files that belong to the member's own history and never move on exploit
# (reference pbt_cluster.py:174,179).
KEEP_NAMES = ("learning_curve.csv", "theta.csv")
KEEP_PREFIXES = ("events.out", ".nfs")

_NP_DTYPES = {
    "float32": np.float32, "float64": np.float64, "float16": np.float16,
    "int32": np.int32, "int64": np.int64, "uint8": np.uint8, "uint16": np.uint16,
    "bfloat16": np.uint16,  # stored as raw 16-bit words
}


def _is_history_file(name: str) -> bool:
    return name in KEEP_NAMES or any(name.startswith(p) for p in KEEP_PREFIXES)


def copy_member_files(src_dir: str, dst_dir: str) -> int:
    """File-level exploit: make ``dst_dir`` a copy of ``src_dir``'s checkpoint.

    Removes every plain file of the destination except its own history files,
    then copies every plain file of the source with the same exclusions;
    sub-directories are skipped.  Returns the number of bytes copied.
    """
    if os.path.abspath(src_dir) == os.path.abspath(dst_dir):
        return 0
    os.makedirs(dst_dir, exist_ok=True)
    for name in os.listdir(dst_dir):
        path = os.path.join(dst_dir, name)
        if not os.path.isdir(path) and not _is_history_file(name):
            try:
                os.remove(path)
            except FileNotFoundError:
                pass
    copied = 0
    if not os.path.isdir(src_dir):
        return 0
    for name in os.listdir(src_dir):
        path = os.path.join(src_dir, name)
        if not os.path.isdir(path) and not _is_history_file(name):
            shutil.copy(path, os.path.join(dst_dir, name))
            copied += os.path.getsize(path)
    return copied


def _align(n: int) -> int:
    return (n + ALIGN - 1) // ALIGN * ALIGN


def _to_numpy(t: Any) -> Tuple[np.ndarray, str]:
    """Accept numpy arrays or torch tensors; returns (array, dtype-name)."""
    if isinstance(t, np.ndarray):
        return np.ascontiguousarray(t), str(t.dtype)
    import torch  # local: keep the module importable without torch
    if isinstance(t, torch.Tensor):
        t = t.detach()
        if t.dtype == torch.bfloat16:
            return t.cpu().contiguous().view(torch.uint16).numpy(), "bfloat16"
        return t.cpu().contiguous().numpy(), str(t.dtype).replace("torch.", "")
    arr = np.asarray(t)
    return np.ascontiguousarray(arr), str(arr.dtype)


def save_checkpoint(save_dir: str, tensors: Mapping[str, Any], *, step: int,
                    optimizer: str = "", meta: Optional[Dict[str, Any]] = None) -> str:
    """Write ``tensors`` as one flat payload; atomic via rename."""
    os.makedirs(save_dir, exist_ok=True)
    entries: Dict[str, Dict[str, Any]] = {}
    offset = 0
    blobs = []
    for name, t in tensors.items():
        arr, dtype = _to_numpy(t)
        entries[name] = {"offset": offset, "shape": list(arr.shape), "dtype": dtype,
                         "nbytes": int(arr.nbytes)}
        blobs.append((offset, arr))
        offset = _align(offset + arr.nbytes)
    index = {"version": FORMAT_VERSION, "step": int(step), "optimizer": optimizer,
             "total_bytes": offset, "entries": entries, "meta": meta or {}}
    data_path = os.path.join(save_dir, CKPT_PREFIX + ".data")
    fd, tmp = tempfile.mkstemp(dir=save_dir, prefix=".tmp_ckpt")
    with os.fdopen(fd, "wb") as f:
        for off, arr in blobs:
            f.seek(off)
            f.write(arr.tobytes())
        f.truncate(offset)
    os.replace(tmp, data_path)
    fd, tmp = tempfile.mkstemp(dir=save_dir, prefix=".tmp_idx")
    with os.fdopen(fd, "w") as f:
        json.dump(index, f, indent=1, sort_keys=True)
    os.replace(tmp, os.path.join(save_dir, CKPT_PREFIX + ".index"))
    with open(os.path.join(save_dir, POINTER_FILE), "w") as f:
        f.write(f'model_checkpoint_path: "{CKPT_PREFIX}"\n'
                f'all_model_checkpoint_paths: "{CKPT_PREFIX}"\n')
    return data_path


def latest_checkpoint(save_dir: str) -> Optional[str]:
    pointer = os.path.join(save_dir, POINTER_FILE)
    if not os.path.isfile(pointer):
        return None
    with open(pointer) as f:
        for line in f:
            if line.startswith("model_checkpoint_path:"):
                prefix = line.split(":", 1)[1].strip().strip('"')
                if os.path.isfile(os.path.join(save_dir, prefix + ".index")):
                    return os.path.join(save_dir, prefix)
    return None


def load_checkpoint(save_dir: str) -> Optional[Tuple[Dict[str, np.ndarray], Dict[str, Any]]]:
    """Returns ``(tensors, index)`` or ``None`` when the directory has no checkpoint.

    bfloat16 entries come back as uint16 arrays (reinterpret with
    ``torch.from_numpy(a).view(torch.bfloat16)``).
    """
    prefix = latest_checkpoint(save_dir)
    if prefix is None:
        return None
    with open(prefix + ".index") as f:
        index = json.load(f)
    if index.get("version") != FORMAT_VERSION:
        raise ValueError(f"unsupported checkpoint version {index.get('version')}")
    raw = np.fromfile(prefix + ".data", dtype=np.uint8)
    out: Dict[str, np.ndarray] = {}
    for name, e in sorted(index["entries"].items(), key=lambda kv: kv[1]["offset"]):   # write order
        dt = np.dtype(_NP_DTYPES[e["dtype"]])
        seg = raw[e["offset"]: e["offset"] + e["nbytes"]]
        out[name] = seg.view(dt).reshape(e["shape"]).copy()
    return out, index


# ----------------------------------------------------------------------------------------------
# Asynchronous save of a native engine's state arena (csrc/host_io.cpp)
# ----------------------------------------------------------------------------------------------
class AsyncCheckpointWriter:
    """Round-end checkpoints that do not stall training: one async device->host copy of the member's state
    region into a pinned staging slot, then a job for the native writer threads (data file, index, pointer, each
    via temp file + rename).  Same on-disk format as ``save_checkpoint``; ``wait`` / ``wait_all`` are the
    barriers restore, the file-exploit path and shutdown use.  (Reference: synchronous ``saver.save`` after
    every train call, mnist_model.py / resnet_run_loop.py.)"""

    def __init__(self, threads: int = 4, staging_slots: int = 4):
        import threading
        from .ops import native
        self._lib = native.lib()
        self._h = self._lib.dtf_io_create(int(threads))
        self._slots = [{"buf": None, "ticket": 0} for _ in range(max(1, staging_slots))]
        self._next = 0
        self._lock = threading.Lock()

    def save_state(self, save_dir: str, state, views: Mapping[str, Any], *, step: int, optimizer: str = "",
                   meta: Optional[Dict[str, Any]] = None) -> int:
        """``state``: contiguous uint8 CUDA tensor; ``views``: name -> tensor view inside it."""
        import torch
        os.makedirs(save_dir, exist_ok=True)
        base, total = state.data_ptr(), state.numel()
        entries: Dict[str, Dict[str, Any]] = {}
        for name, t in views.items():
            nbytes = t.numel() * t.element_size()
            off = t.data_ptr() - base if nbytes else 0       # empty views (e.g. no BN statistics) carry no pointer
            if off < 0 or off + nbytes > total or not t.is_contiguous():
                raise ValueError(f"{name} is not a contiguous view of the state region")
            entries[name] = {"offset": int(off), "shape": list(t.shape), "dtype": str(t.dtype).replace("torch.", ""),
                             "nbytes": int(nbytes)}
        index = {"version": FORMAT_VERSION, "step": int(step), "optimizer": optimizer, "total_bytes": int(total),
                 "entries": entries, "meta": meta or {}}
        with self._lock:
            slot = self._slots[self._next]
            self._next = (self._next + 1) % len(self._slots)
            if slot["ticket"]:
                self._lib.dtf_io_wait(self._h, slot["ticket"])          # the slot's previous job is on disk
            if slot["buf"] is None or slot["buf"].numel() < total:
                slot["buf"] = torch.empty(total, dtype=torch.uint8, pin_memory=state.is_cuda)
            host = slot["buf"][:total]
            host.copy_(state, non_blocking=True)
            pointer = (f'model_checkpoint_path: "{CKPT_PREFIX}"\n'
                       f'all_model_checkpoint_paths: "{CKPT_PREFIX}"\n')
            ticket = self._lib.dtf_io_submit(
                self._h, os.path.join(save_dir, CKPT_PREFIX + ".data").encode(), host.data_ptr(), total,
                torch.cuda.current_stream(state.device).cuda_stream if state.is_cuda else None,
                (state.device.index or 0) if state.is_cuda else -1,
                os.path.join(save_dir, CKPT_PREFIX + ".index").encode(),
                json.dumps(index, indent=1, sort_keys=True).encode(),
                os.path.join(save_dir, POINTER_FILE).encode(), pointer.encode())
            if ticket <= 0:
                raise RuntimeError(f"dtf_io_submit failed: {ticket}")
            slot["ticket"] = ticket
            return int(ticket)

    def wait(self, ticket: int) -> None:
        if ticket and self._lib.dtf_io_wait(self._h, int(ticket)) != 0:
            raise RuntimeError("checkpoint writer reported a failed write (see stderr)")

    def wait_all(self) -> None:
        if self._lib.dtf_io_wait_all(self._h) != 0:
            raise RuntimeError("checkpoint writer reported a failed write (see stderr)")

    def pending(self) -> int:
        return int(self._lib.dtf_io_pending(self._h))

    def bytes_written(self) -> int:
        return int(self._lib.dtf_io_bytes_written(self._h))

    def close(self) -> None:
        if self._h:
            self._lib.dtf_io_destroy(self._h)
            self._h = None


_WRITER: Optional[AsyncCheckpointWriter] = None


def async_writer() -> AsyncCheckpointWriter:
    """Process-wide writer (created on first use; drained at interpreter exit)."""
    global _WRITER
    if _WRITER is None:
        import atexit
        _WRITER = AsyncCheckpointWriter()
        atexit.register(lambda: _WRITER is not None and _WRITER._h and _WRITER.wait_all())
    return _WRITER
