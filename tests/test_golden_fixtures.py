"""Golden numeric fixtures of the reference (TF 1.8) vs the functional ResNet oracle.

The fixtures live in the read-only reference tree; they are read in place (never
copied) and the test is skipped where the tree is absent (e.g. on the GPU box).
See SURVEY §4.2 for the decoded fixture configurations.
"""
import json
import os
import re

import numpy as np
import pytest
import torch

from distributedtf_b200.models import resnet
from distributedtf_b200.utils.testing.tf_bundle import read_bundle

FIXTURES = "/root/reference/resnet/official/utils/testing/reference_data/resnet"
pytestmark = pytest.mark.skipif(not os.path.isdir(FIXTURES), reason="reference fixtures not mounted")


def _cases():
    if not os.path.isdir(FIXTURES):
        return []
    return sorted(d for d in os.listdir(FIXTURES) if d.startswith("batch-size"))


def _fingerprint(t_nhwc: np.ndarray):
    flat = t_nhwc.reshape(-1)
    return list(t_nhwc.shape) + [float(flat[0]), float(flat[-1]), float(flat.astype(np.float64).sum())]


def _check(expected, got, n_elems_out):
    exp_in, exp_out = expected[:7], expected[7:]
    got_in, got_out = got[:7], got[7:]
    assert got_in[:4] == exp_in[:4] and got_out[:4] == exp_out[:4]
    np.testing.assert_allclose(got_in[4:], exp_in[4:], rtol=1e-5)
    np.testing.assert_allclose(got_out[4:6], exp_out[4:6], rtol=2e-4, atol=2e-5)
    # sums are fp32 accumulations in TF: compare with an absolute tolerance scaled by size
    assert abs(got_out[6] - exp_out[6]) < max(0.1, 2e-4 * abs(exp_out[6])), (got_out[6], exp_out[6])


def test_batch_norm_fixture():
    b = read_bundle(os.path.join(FIXTURES, "batch_norm", "model.ckpt"))
    expected = json.load(open(os.path.join(FIXTURES, "batch_norm", "results.json")))
    x = torch.from_numpy(b["input_tensor"]).permute(0, 3, 1, 2)
    y = resnet.batch_norm(x, torch.from_numpy(b["batch_normalization/gamma"]),
                          torch.from_numpy(b["batch_normalization/beta"]),
                          torch.from_numpy(b["batch_normalization/moving_mean"]).clone(),
                          torch.from_numpy(b["batch_normalization/moving_variance"]).clone(), training=True)
    got = _fingerprint(b["input_tensor"]) + _fingerprint(y.permute(0, 2, 3, 1).contiguous().numpy())
    _check(expected, got, y.numel())


@pytest.mark.parametrize("case", _cases())
def test_block_fixture(case):
    m = re.match(r"batch-size-(\d+)_(building|bottleneck)(_projection)?_version-(\d)_width-(\d+)_channels-(\d+)", case)
    bottleneck, projection, version = m.group(2) == "bottleneck", bool(m.group(3)), int(m.group(4))
    b = read_bundle(os.path.join(FIXTURES, case, "model.ckpt"))
    expected = json.load(open(os.path.join(FIXTURES, case, "results.json")))
    params = {}
    for name, arr in b.items():
        mm = re.match(r"(conv2d|batch_normalization)(?:_(\d+))?/(\w+)", name)
        if not mm:
            continue
        idx = int(mm.group(2) or 0)
        t = torch.from_numpy(arr)
        if mm.group(1) == "conv2d":
            params[f"conv{idx}/kernel"] = t.permute(3, 0, 1, 2).contiguous()      # HWIO -> OHWI
        else:
            params[f"bn{idx}/{mm.group(3)}"] = t.clone()
    if projection:
        stride, filters = 2, (2 if bottleneck else 8)
    else:
        stride, filters = 1, (1 if bottleneck else 4)
    x = torch.from_numpy(b["input_tensor"]).permute(0, 3, 1, 2)
    y = resnet.run_block(version, bottleneck, projection, filters, stride, params, x, training=True)
    got = _fingerprint(b["input_tensor"]) + _fingerprint(y.permute(0, 2, 3, 1).contiguous().numpy())
    _check(expected, got, y.numel())
    # the parameter walk names/creates exactly the variables stored in the fixture
    specs = resnet.block_param_specs(version, bottleneck, projection, 4, filters)
    assert {s.name for s in specs} == set(params)
    assert all(tuple(params[s.name].shape) == s.shape for s in specs)
