"""HP sampler support + distribution (reference constants.py:45-100)."""
import collections
import random

from distributedtf_b200 import hpspace


def test_sample_shape_and_support():
    rng = random.Random(0)
    seen_opt = collections.Counter()
    for _ in range(3000):
        hp = hpspace.generate_random_hparam(rng)
        assert set(hp) == {"opt_case", "decay_steps", "decay_rate", "weight_decay",
                           "regularizer", "initializer", "batch_size"}
        assert hpspace.validate_hparam(hp) == [], hp
        assert isinstance(hp["batch_size"], int) and 65 <= hp["batch_size"] <= 255
        oc = hp["opt_case"]
        assert oc["lr"] in hpspace.get_hp_range_definition()["lr"][oc["optimizer"]]
        assert hp["decay_steps"] in range(0, 101, 10)
        seen_opt[oc["optimizer"]] += 1
    assert set(seen_opt) == set(hpspace.OPTIMIZERS)
    # uniform choice over 6 optimizers: each ~500 of 3000
    assert all(350 < c < 650 for c in seen_opt.values()), seen_opt


def test_branch_keys():
    rng = random.Random(1)
    for _ in range(500):
        oc = hpspace.generate_random_hparam(rng)["opt_case"]
        keys = set(oc)
        if oc["optimizer"] == "Momentum":
            assert keys == {"optimizer", "lr", "momentum"}
        elif oc["optimizer"] == "RMSProp":
            assert keys == {"optimizer", "lr", "momentum", "grad_decay"}
        else:
            assert keys == {"optimizer", "lr"}


def test_batch_size_covers_range_ends():
    rng = random.Random(2)
    sizes = {hpspace.generate_random_hparam(rng)["batch_size"] for _ in range(20000)}
    assert min(sizes) == 65 and max(sizes) == 255


def test_seeded_sampling_is_reproducible():
    a = [hpspace.generate_random_hparam(random.Random(7)) for _ in range(3)]
    b = [hpspace.generate_random_hparam(random.Random(7)) for _ in range(3)]
    assert a == b


def test_samples_do_not_alias():
    rng = random.Random(3)
    a, b = hpspace.generate_random_hparam(rng), hpspace.generate_random_hparam(rng)
    a["opt_case"]["lr"] = -1
    assert b["opt_case"]["lr"] != -1
    assert -1 not in hpspace.get_hp_range_definition()["lr"][b["opt_case"]["optimizer"]]


def test_default_cifar_point_is_inside_the_space():
    assert hpspace.validate_hparam(hpspace.default_cifar_hparam()) == []
