"""Datasets and batch iteration for the MNIST / CIFAR-10 members.

Parity: reference ``resnet/cifar10_main.py:54-140`` (3073-byte records, CHW->HWC,
pad to 40x40 + random 32x32 crop + random horizontal flip for training,
per-image standardisation always), ``resnet_run_loop.py:45-105`` (shuffle over the
whole training set, repeat, batch, prefetch), ``mnist_model.py:131-138,153-171``
(idx-gz files, float32 pixels 0..255 *unnormalised*, shuffled numpy batches) and
``official/utils/misc/model_helpers.py:59-86`` (synthetic data = constant zeros
with label 0).

B200-first: the whole uint8 dataset is uploaded once and lives on the GPU (CIFAR
train = 150 MB, MNIST = 47 MB, nothing next to 180 GB); a batch is an index
gather + augmentation kernel on the device, so the input pipeline never touches
the host in steady state.  ``synthetic="random"`` (default for benchmarks) draws
fixed random uint8 images/labels instead of the reference's zeros, because zero
images make every BatchNorm variance zero and the work unrepresentative.
"""
from __future__ import annotations

import gzip
import os
from dataclasses import dataclass
from typing import Iterator, Optional, Tuple

import numpy as np
import torch

CIFAR_TRAIN, CIFAR_EVAL = 50000, 10000
MNIST_TRAIN, MNIST_EVAL = 60000, 10000
_CIFAR_RECORD = 1 + 3 * 32 * 32


@dataclass
class Dataset:
    """uint8 images ``[N, H, W, C]`` (NHWC) + int32 labels, resident on ``device``."""
    images: torch.Tensor
    labels: torch.Tensor
    name: str = ""

    def __len__(self) -> int:
        return int(self.images.shape[0])

    def to(self, device) -> "Dataset":
        return Dataset(self.images.to(device), self.labels.to(device), self.name)


def synthetic_dataset(n: int, h: int, w: int, c: int, num_classes: int, mode: str = "random",
                      seed: int = 0, device="cpu") -> Dataset:
    if mode == "zeros":   # the reference's synthetic data: one constant element repeated (model_helpers.py:59-86)
        from ..misc.model_helpers import generate_synthetic_data
        images, labels = next(generate_synthetic_data([h, w, c], 0, torch.uint8, [], 0, torch.int32,
                                                      batch_size=n, device=device))
        return Dataset(images, labels, "synthetic-zeros")
    g = torch.Generator().manual_seed(seed)
    images = torch.randint(0, 256, (n, h, w, c), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, num_classes, (n,), dtype=torch.int32, generator=g)
    return Dataset(images.to(device), labels.to(device), "synthetic-random")


def load_cifar10(data_dir: str, train: bool, device="cpu") -> Dataset:
    root = os.path.join(data_dir, "cifar-10-batches-bin")
    if not os.path.isdir(root):
        raise FileNotFoundError(f"{root} not found (expected the CIFAR-10 binary version)")
    names = [f"data_batch_{i}.bin" for i in range(1, 6)] if train else ["test_batch.bin"]
    paths = [os.path.join(root, n) for n in names]
    native = _native_lib()
    if native is not None:                    # csrc/host_data.cpp: threaded record decode straight into the tensors
        sizes = [os.path.getsize(p) for p in paths]
        if any(sz % _CIFAR_RECORD for sz in sizes):
            raise ValueError("CIFAR-10 binary file size is not a multiple of the 3073-byte record")
        n = sum(sizes) // _CIFAR_RECORD
        raw = torch.empty(sum(sizes), dtype=torch.uint8)
        off = 0
        for p, sz in zip(paths, sizes):
            got = native.dtf_read_file(p.encode(), 0, raw.data_ptr() + off, sz)
            if got != sz:
                raise IOError(f"short read on {p}: {got} of {sz} bytes")
            off += sz
        images = torch.empty(n, 32, 32, 3, dtype=torch.uint8)
        labels = torch.empty(n, dtype=torch.int32)
        rc = native.dtf_cifar_decode(raw.data_ptr(), n, images.data_ptr(), labels.data_ptr(), min(8, os.cpu_count() or 1))
        if rc != 0:
            raise RuntimeError(f"dtf_cifar_decode failed: {rc}")
        return Dataset(images.to(device), labels.to(device), "CIFAR-10")
    raw = np.concatenate([np.fromfile(p, dtype=np.uint8) for p in paths])
    rec = raw.reshape(-1, _CIFAR_RECORD)
    labels = torch.from_numpy(rec[:, 0].astype(np.int32))
    images = torch.from_numpy(rec[:, 1:].reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1).copy())
    return Dataset(images.to(device), labels.to(device), "CIFAR-10")


def _native_lib():
    """The native host runtime when the library is built (readers fall back to numpy otherwise)."""
    try:
        from ...ops import native
        return native.lib() if native.available() else None
    except Exception:
        return None


def load_mnist(data_dir: str, train: bool, device="cpu") -> Dataset:
    stem = "train" if train else "t10k"
    with gzip.open(os.path.join(data_dir, f"{stem}-images-idx3-ubyte.gz"), "rb") as f:
        images = np.frombuffer(f.read(), np.uint8, offset=16).reshape(-1, 28, 28, 1).copy()
    with gzip.open(os.path.join(data_dir, f"{stem}-labels-idx1-ubyte.gz"), "rb") as f:
        labels = np.frombuffer(f.read(), np.uint8, offset=8).astype(np.int32)
    return Dataset(torch.from_numpy(images).to(device), torch.from_numpy(labels).to(device), "MNIST")


def cifar_preprocess_torch(images_u8: torch.Tensor, training: bool,
                           gen: Optional[torch.Generator] = None) -> torch.Tensor:
    """Reference preprocessing with torch ops (oracle for the augmentation kernel).
    uint8 NHWC in, float32 NHWC out."""
    x = images_u8.float()
    n, h, w, c = x.shape
    if training:
        dev = x.device
        padded = torch.zeros(n, h + 8, w + 8, c, device=dev)
        padded[:, 4:4 + h, 4:4 + w] = x
        oy = torch.randint(0, 9, (n,), generator=gen, device=gen.device if gen else "cpu").to(dev)
        ox = torch.randint(0, 9, (n,), generator=gen, device=gen.device if gen else "cpu").to(dev)
        flip = (torch.rand(n, generator=gen, device=gen.device if gen else "cpu") < 0.5).to(dev)
        rows = (oy[:, None] + torch.arange(h, device=dev)[None])            # [n, h]
        cols = (ox[:, None] + torch.arange(w, device=dev)[None])            # [n, w]
        cols = torch.where(flip[:, None], cols.flip(1), cols)
        x = padded[torch.arange(n, device=dev)[:, None, None], rows[:, :, None], cols[:, None, :]]
    # tf.image.per_image_standardization: (x - mean) / max(std, 1/sqrt(num_elements))
    mean = x.mean(dim=(1, 2, 3), keepdim=True)
    std = x.std(dim=(1, 2, 3), keepdim=True, unbiased=False)
    return (x - mean) / torch.clamp(std, min=1.0 / float(h * w * c) ** 0.5)


class BatchIterator:
    """Endless shuffled batches of indices over a resident dataset (drop remainder)."""

    def __init__(self, n: int, batch_size: int, shuffle: bool, seed: int = 0, device="cpu"):
        self.n, self.bs, self.shuffle = n, batch_size, shuffle
        self.device = device
        self._gen = torch.Generator().manual_seed(seed)
        self._perm = None
        self._pos = 0

    def set_batch_size(self, bs: int) -> None:
        self.bs = bs

    def next_indices(self) -> torch.Tensor:
        if self._perm is None or self._pos + self.bs > self.n:
            self._perm = (torch.randperm(self.n, generator=self._gen) if self.shuffle
                          else torch.arange(self.n)).to(self.device)
            self._pos = 0
        idx = self._perm[self._pos: self._pos + self.bs]
        self._pos += self.bs
        return idx

    def epoch_batches(self) -> Iterator[torch.Tensor]:
        """One pass, in order, *including* the final short batch (evaluation)."""
        order = torch.arange(self.n, device=self.device)
        for s in range(0, self.n, self.bs):
            yield order[s: s + self.bs]
