"""The device hyper-parameter block must hold every LR schedule explore can produce."""
import ctypes as C
import types

import pytest

from distributedtf_b200.models.native_base import NativeEngineBase
from distributedtf_b200.ops import native, reference


def _stub():
    s = types.SimpleNamespace(uses_regularizer=True)
    s.lr_schedule = lambda hp: reference.cifar_lr_schedule(hp)
    return s


@pytest.mark.parametrize("decay_steps", list(range(0, 12)) + [20, 50, 99, 100])
def test_every_decay_steps_value_fits_the_device_block(decay_steps):
    """explore perturbs decay_steps as an int in [0, 100] (reference model_base.py:54-68): small values give
    up to 99 boundaries and tf.train.piecewise_constant takes any number, so the block must too."""
    hp = {"opt_case": {"optimizer": "Momentum", "lr": 0.1, "momentum": 0.9}, "decay_steps": decay_steps,
          "decay_rate": 0.5, "weight_decay": 2e-4, "regularizer": "l2_regularizer", "batch_size": 65}
    h = NativeEngineBase.hparams_struct(_stub(), hp)
    bounds, values = reference.cifar_lr_schedule(hp)
    assert h.n_boundaries == len(bounds) <= native.MAX_LR_BOUNDARIES
    assert [h.boundaries[i] for i in range(h.n_boundaries)] == [int(b) for b in bounds]
    assert h.lr_values[len(values) - 1] == pytest.approx(values[-1], rel=1e-6)


def test_block_layout_matches_the_header():
    # 8 scalars + 100 int64 boundaries + 101 lr values + 3 reserved floats, a multiple of 16 bytes for the
    # vector copy in the exploit kernel
    assert native.MAX_LR_BOUNDARIES == 100
    assert C.sizeof(native.Hparams) == 32 + 800 + 404 + 12
    assert C.sizeof(native.Hparams) % 16 == 0
