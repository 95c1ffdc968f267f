"""Hyper-parameter ranges and the initial-population sampler.

Parity: reference ``constants.py:14-43`` (range table), ``:45-94`` (search space
built with hyperopt ``hp.choice / hp.uniform / hp.randint``) and ``:96-100``
(``generate_random_hparam``).  hyperopt is not available on the target image, so
the space is expressed with three tiny node types that have the same sampling
semantics as ``hyperopt.pyll.stochastic.sample``:

* ``Choice(options)``   — uniform over the options; an option may itself be a
  dict containing further nodes (the nested ``opt_case`` branch);
* ``Uniform(lo, hi)``   — float uniform on [lo, hi);
* ``RandInt(n) + k``    — integer uniform on [k, k + n).

The sampled dictionary has exactly the reference's shape::

    {'opt_case': {'optimizer': str, 'lr': float[, 'momentum'][, 'grad_decay']},
     'decay_steps': int, 'decay_rate': float, 'weight_decay': float,
     'regularizer': str, 'initializer': str, 'batch_size': int in [65, 255]}
"""
from __future__ import annotations

import copy
import random
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence

OPTIMIZERS = ("Adadelta", "Adagrad", "Momentum", "Adam", "RMSProp", "gd")
REGULARIZERS = ("l1_regularizer", "l2_regularizer", "l1_l2_regularizer", "None")
INITIALIZERS = ("glorot_normal", "orthogonal", "he_init", "None")

BATCH_SIZE_OFFSET = 65          # batch_size = randint(191) + 65  ->  [65, 255]
BATCH_SIZE_SPAN = 191


def get_hp_range_definition() -> Dict[str, Any]:
    """Range table consulted by both the sampler and ``explore.perturb``.

    For list-valued entries the first and last element are the clipping bounds
    used by explore (reference ``model_base.py:70-79``).
    """
    return {
        "h_0": [0.0, 1.0],
        "h_1": [0.0, 1.0],
        "optimizer_list": list(OPTIMIZERS),
        "lr": {
            "Adadelta": [0.1, 0.5, 1.0, 1.5, 2.0, 2.5, 3.0, 3.5, 4.0, 4.5, 5.0],
            "Adagrad": [1e-3, 1e-2, 1e-1, 0.5, 1.0],
            "Momentum": [1e-3, 1e-2, 1e-1, 0.5, 1.0],
            "Adam": [1e-4, 1e-3, 1e-2, 1e-1],
            "RMSProp": [1e-5, 1e-4, 1e-3],
            "gd": [1e-2, 1e-1, 0.5, 1.0, 1.5, 2.0, 2.5, 3.0, 3.5, 4.0, 4.5, 5.0],
        },
        "momentum": [0.00, 0.9],
        "grad_decay": [0.00, 0.9],
        "decay_steps": [0, 10, 20, 30, 40, 50, 60, 70, 80, 90, 100],
        "decay_rate": [0.1, 1.0],
        "weight_decay": [1e-8, 1e-2],
        "regularizer": list(REGULARIZERS),
        "initializer": list(INITIALIZERS),
        "batch_size": [BATCH_SIZE_SPAN],
    }


# --------------------------------------------------------------------------- nodes
class Node:
    def sample(self, rng: random.Random) -> Any:  # pragma: no cover - interface
        raise NotImplementedError


@dataclass(frozen=True)
class Choice(Node):
    label: str
    options: Sequence[Any]

    def sample(self, rng: random.Random) -> Any:
        return _materialise(self.options[rng.randrange(len(self.options))], rng)


@dataclass(frozen=True)
class Uniform(Node):
    label: str
    low: float
    high: float

    def sample(self, rng: random.Random) -> float:
        return rng.uniform(self.low, self.high)


@dataclass(frozen=True)
class RandInt(Node):
    label: str
    upper: int
    offset: int = 0

    def sample(self, rng: random.Random) -> int:
        return rng.randrange(self.upper) + self.offset

    def __add__(self, k: int) -> "RandInt":
        return RandInt(self.label, self.upper, self.offset + int(k))


def _materialise(obj: Any, rng: random.Random) -> Any:
    if isinstance(obj, Node):
        return obj.sample(rng)
    if isinstance(obj, dict):
        return {k: _materialise(v, rng) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_materialise(v, rng) for v in obj)
    return copy.deepcopy(obj)


def sample_space(space: Any, rng: Optional[random.Random] = None) -> Any:
    """Draw one point of a (nested) space — ``hyperopt.pyll.stochastic.sample``."""
    return _materialise(space, rng or random)


# --------------------------------------------------------------------------- space
def load_hp_space() -> Dict[str, Any]:
    r = get_hp_range_definition()

    def branch(name: str, momentum: bool = False, grad_decay: bool = False) -> dict:
        b: Dict[str, Any] = {"optimizer": name, "lr": Choice("lr", r["lr"][name])}
        if grad_decay:
            b["grad_decay"] = Uniform("grad_decay", r["grad_decay"][0], r["grad_decay"][1])
        if momentum:
            b["momentum"] = Uniform("momentum", r["momentum"][0], r["momentum"][1])
        return b

    return {
        "opt_case": Choice("opt_case", [
            branch("Adadelta"),
            branch("Adagrad"),
            branch("Momentum", momentum=True),
            branch("Adam"),
            branch("RMSProp", momentum=True, grad_decay=True),
            branch("gd"),
        ]),
        "decay_steps": Choice("decay_steps", r["decay_steps"]),
        "decay_rate": Uniform("decay_rate", r["decay_rate"][0], r["decay_rate"][1]),
        "weight_decay": Uniform("weight_decay", r["weight_decay"][0], r["weight_decay"][1]),
        "regularizer": Choice("regularizer", r["regularizer"]),
        "initializer": Choice("initializer", r["initializer"]),
        "batch_size": RandInt("batch_size", r["batch_size"][0]) + BATCH_SIZE_OFFSET,
    }


def generate_random_hparam(rng: Optional[random.Random] = None) -> Dict[str, Any]:
    sample = sample_space(load_hp_space(), rng)
    sample["batch_size"] = int(sample["batch_size"])
    return sample


def default_cifar_hparam() -> Dict[str, Any]:
    """The hand-tuned CIFAR point left commented in ``cifar10_model.py:15-22``;
    used as the fixed configuration of the throughput benchmark."""
    return {
        "opt_case": {"optimizer": "Momentum", "lr": 0.1, "momentum": 0.9},
        "decay_steps": 20,
        "decay_rate": 0.1,
        "weight_decay": 2e-4,
        "regularizer": "l2_regularizer",
        "initializer": "he_init",
        "batch_size": 128,
    }


def validate_hparam(hp: Dict[str, Any]) -> List[str]:
    """Return a list of problems (empty when ``hp`` lies inside the space)."""
    r = get_hp_range_definition()
    bad: List[str] = []
    oc = hp.get("opt_case")
    if not isinstance(oc, dict) or oc.get("optimizer") not in OPTIMIZERS:
        return ["opt_case.optimizer missing or unknown"]
    name = oc["optimizer"]
    lo, hi = r["lr"][name][0], r["lr"][name][-1]
    if not (lo <= oc["lr"] <= hi):
        bad.append(f"lr {oc['lr']} outside [{lo},{hi}] for {name}")
    want = {"optimizer", "lr"}
    if name in ("Momentum", "RMSProp"):
        want.add("momentum")
    if name == "RMSProp":
        want.add("grad_decay")
    if set(oc) != want:
        bad.append(f"opt_case keys {sorted(oc)} != {sorted(want)}")
    for k in ("momentum", "grad_decay"):
        if k in oc and not (r[k][0] <= oc[k] <= r[k][-1]):
            bad.append(f"{k} {oc[k]} out of range")
    if "decay_steps" in hp and not (0 <= hp["decay_steps"] <= 100):
        bad.append("decay_steps out of range")
    for k in ("decay_rate", "weight_decay"):
        if k in hp and not (r[k][0] <= hp[k] <= r[k][-1]):
            bad.append(f"{k} {hp[k]} out of range")
    if "regularizer" in hp and hp["regularizer"] not in REGULARIZERS:
        bad.append("regularizer unknown")
    if "initializer" in hp and hp["initializer"] not in INITIALIZERS:
        bad.append("initializer unknown")
    if "batch_size" in hp and not (BATCH_SIZE_OFFSET <= hp["batch_size"] <= BATCH_SIZE_OFFSET + BATCH_SIZE_SPAN):
        bad.append("batch_size out of range")
    return bad
