"""Property tests for explore/perturb (reference model_base.py:30-104)."""
import copy
import random

import pytest
from hypothesis import given, settings, strategies as st

from distributedtf_b200 import explore, hpspace
from distributedtf_b200.model_base import ModelBase


@pytest.mark.parametrize("limit,digits", [(1e-08, 8), (0.0001, 4), (0.0, 1), (0.1, 1),
                                          (1e-05, 5), (0.001, 3), (0.01, 2)])
def test_digits_follow_the_text_of_the_minimum(limit, digits):
    assert explore.digits_of(limit) == digits


@settings(max_examples=300, deadline=None)
@given(seed=st.integers(0, 2 ** 31), val=st.floats(0.1, 1.0))
def test_float_stays_in_range_and_is_rounded(seed, val):
    rng = random.Random(seed)
    out = explore.perturb_float(val, 0.1, 1.0, rng)
    assert 0.1 - 1e-12 <= out <= 1.0 + 1e-12
    assert 0.8 * val - 0.06 <= out <= 1.2 * val + 0.06     # within factors up to rounding
    # 1 digit, or 2 when the lower clip was hit
    assert round(out, 2) == out


def test_lower_clip_adds_one_digit():
    rng = random.Random(0)
    outs = {explore.perturb_float(1e-8, 1e-8, 1e-2, rng) for _ in range(50)}
    assert all(round(o, 9) == o for o in outs)
    assert any(round(o, 8) != o for o in outs)   # the 9th digit is actually used


@settings(max_examples=300, deadline=None)
@given(seed=st.integers(0, 2 ** 31), val=st.integers(65, 256))
def test_batch_size_int_rule(seed, val):
    hp = {"batch_size": val}
    explore.perturb_hparams(hp, random.Random(seed))
    assert isinstance(hp["batch_size"], int)
    assert 65 <= hp["batch_size"] <= 256
    assert int(0.8 * val) - 1 <= hp["batch_size"] <= int(1.2 * val) + 1


def test_int_rule_degenerate_and_bounds():
    assert explore.perturb_int(0, 0, 100, random.Random(0)) == 0
    assert all(80 <= explore.perturb_int(100, 0, 100, random.Random(s)) <= 100 for s in range(50))


@settings(max_examples=200, deadline=None)
@given(seed=st.integers(0, 2 ** 31))
def test_full_hparam_perturbation_invariants(seed):
    rng = random.Random(seed)
    hp = hpspace.generate_random_hparam(rng)
    before = copy.deepcopy(hp)
    explore.perturb_hparams(hp, rng)
    assert hp["opt_case"]["optimizer"] == before["opt_case"]["optimizer"]     # optimizer frozen
    assert set(hp["opt_case"]) == set(before["opt_case"])
    assert hp["regularizer"] == before["regularizer"]                         # frozen strings
    assert hp["initializer"] == before["initializer"]
    assert hpspace.validate_hparam({**hp, "batch_size": min(hp["batch_size"], 256)}) == [] or \
        hp["batch_size"] == 256
    r = hpspace.get_hp_range_definition()
    lo, hi = r["lr"][hp["opt_case"]["optimizer"]][0], r["lr"][hp["opt_case"]["optimizer"]][-1]
    assert lo <= hp["opt_case"]["lr"] <= hi
    assert 0 <= hp["decay_steps"] <= 100 and isinstance(hp["decay_steps"], int)
    assert 0.1 <= hp["decay_rate"] <= 1.0
    assert 1e-8 <= hp["weight_decay"] <= 1e-2


def test_toy_h_values_are_perturbed_as_floats():
    hp = {"h_0": 1.0, "h_1": 0.0}
    explore.perturb_hparams(hp, random.Random(0))
    assert 0.8 <= hp["h_0"] <= 1.0 and hp["h_1"] == 0.0


def test_model_base_contract(tmp_path):
    hp = hpspace.generate_random_hparam(random.Random(0))
    m = ModelBase(3, hp, str(tmp_path / "model_"))
    assert m.get_values() == [3, 0.0, hp]
    assert (m.epoches_trained, m.need_explore, m._perturb_factors) == (0, False, [0.8, 1.2])
    with pytest.raises(NotImplementedError):
        m.train(1, 1)
    winner_hp = hpspace.generate_random_hparam(random.Random(1))
    m.set_values([3, 0.9, winner_hp])
    assert m.hparams == winner_hp and m.hparams is not winner_hp
    m.seed(0)
    m.perturb_hparams()          # must not throw (reference test_mnist_deep_model.py:9-27)
    assert m.save_dir.endswith("model_3")
