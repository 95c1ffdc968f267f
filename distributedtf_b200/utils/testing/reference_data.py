"""Golden-data harness.

Parity: reference ``official/utils/testing/reference_data.py:62-334`` — a test case stores, for a named
computation, the input/output fingerprints ``[shape..., first, last, sum]`` (``common_tensor_properties``)
next to the tensors that reproduce it, and later runs compare against the stored numbers.  TensorFlow
graphs/checkpoints become: a callable + its input tensors saved in the ``ckpt.py`` format, results in
``results.json``.  ``read_reference_fixture`` additionally reads the *reference's own* TF-1.8 fixtures
(through ``tf_bundle``) so they serve as a TensorFlow-free oracle.
"""
from __future__ import annotations

import hashlib
import json
import os
from typing import Callable, Dict, List, Sequence

import numpy as np
import torch

from ... import ckpt
from .tf_bundle import read_bundle


def name_to_seed(name: str) -> int:
    """Stable 32-bit seed derived from a test name."""
    return int(hashlib.md5(name.encode("utf-8")).hexdigest(), 16) % (2 ** 32 - 1)


def common_tensor_properties(t) -> List[float]:
    """``shape..., first element, last element, sum`` — cheap fingerprint that catches most changes."""
    a = t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
    flat = a.reshape(-1)
    return list(a.shape) + [float(flat[0]), float(flat[-1]), float(flat.astype(np.float64).sum())]


def default_correctness_function(*tensors) -> List[float]:
    out: List[float] = []
    for t in tensors:
        out += common_tensor_properties(t)
    return out


class ReferenceData:
    def __init__(self, data_root: str, test_name: str):
        self.dir = os.path.join(data_root, test_name)

    def save(self, inputs: Dict[str, torch.Tensor], fn: Callable[..., Sequence[torch.Tensor]],
             correctness_function=default_correctness_function) -> List[float]:
        outs = fn(**inputs)
        outs = outs if isinstance(outs, (list, tuple)) else [outs]
        results = correctness_function(*list(inputs.values()), *outs)
        ckpt.save_checkpoint(self.dir, inputs, step=0, meta={"kind": "reference_data"})
        with open(os.path.join(self.dir, "results.json"), "w") as f:
            json.dump(results, f)
        with open(os.path.join(self.dir, "version.json"), "w") as f:
            json.dump({"torch": torch.__version__}, f)
        return results

    def check(self, fn: Callable[..., Sequence[torch.Tensor]], correctness_function=default_correctness_function,
              rtol: float = 1e-5, atol: float = 1e-6) -> None:
        tensors, _ = ckpt.load_checkpoint(self.dir)
        inputs = {k: torch.from_numpy(v) for k, v in tensors.items()}
        outs = fn(**inputs)
        outs = outs if isinstance(outs, (list, tuple)) else [outs]
        got = correctness_function(*list(inputs.values()), *outs)
        with open(os.path.join(self.dir, "results.json")) as f:
            want = json.load(f)
        if len(got) != len(want):
            raise AssertionError(f"fingerprint length changed: {len(got)} vs {len(want)}")
        np.testing.assert_allclose(got, want, rtol=rtol, atol=atol)


def read_reference_fixture(case_dir: str):
    """(variables, expected results) of one of the reference's TF-1.8 golden fixtures."""
    with open(os.path.join(case_dir, "results.json")) as f:
        expected = json.load(f)
    return read_bundle(os.path.join(case_dir, "model.ckpt")), expected
