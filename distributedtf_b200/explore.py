"""Explore step: perturb a member's hyper-parameters in place.

Parity: reference ``model_base.py:30-104``.  Rules:

* float   -> uniform in [0.8 v, 1.2 v], clipped to the range table, rounded to a
  digit count read off the *textual form of the range minimum* (``1e-08`` -> 8,
  ``0.0001`` -> 4, ``0.0`` / ``0.1`` -> 1), one more digit when the lower clip hit;
* int     -> ``randint(floor(0.8 v), ceil(1.2 v))`` clipped; ``batch_size`` is
  clipped to [65, 256];
* str     -> ``regularizer`` / ``initializer`` (and architecture keys) frozen,
  any other string key re-drawn from its list;
* ``opt_case`` -> optimizer *type* frozen (required so optimizer slots stay
  meaningful after exploit); ``lr`` perturbed inside that optimizer's list bounds,
  ``momentum`` for Momentum/RMSProp, ``grad_decay`` for RMSProp.
"""
from __future__ import annotations

import math
import random
from typing import Any, Dict, Optional, Tuple

from .hpspace import BATCH_SIZE_OFFSET, get_hp_range_definition

PERTURB_FACTORS: Tuple[float, float] = (0.8, 1.2)
FROZEN_STRING_KEYS = frozenset({
    "num_filters_1", "kernel_size_1", "kernel_size_2", "activation",
    "initializer", "regularizer",
})


def digits_of(limit_min: float) -> int:
    """Rounding precision implied by how ``limit_min`` prints."""
    text = str(limit_min)
    if "e" in text:
        exponent = int(text.split("e")[1])
        return -exponent if exponent < 0 else exponent
    return text[::-1].find(".")


def perturb_float(val: float, limit_min: float, limit_max: float,
                  rng: Optional[random.Random] = None,
                  factors: Tuple[float, float] = PERTURB_FACTORS) -> float:
    rng = rng or random
    ndigits = digits_of(limit_min)
    lo, hi = val * factors[0], val * factors[1]
    if lo < limit_min:
        lo = limit_min
        ndigits += 1
    if hi > limit_max:
        hi = limit_max
    return round(rng.uniform(lo, hi), ndigits)


def perturb_int(val: int, limit_min: int, limit_max: int,
                rng: Optional[random.Random] = None,
                factors: Tuple[float, float] = PERTURB_FACTORS) -> int:
    rng = rng or random
    if limit_min == limit_max:
        limit_min = 0
    lo = max(int(math.floor(val * factors[0])), limit_min)
    hi = min(int(math.ceil(val * factors[1])), limit_max)
    if lo >= hi:
        return lo
    return rng.randint(lo, hi)


def perturb_hparams(hparams: Dict[str, Any], rng: Optional[random.Random] = None,
                    factors: Tuple[float, float] = PERTURB_FACTORS) -> Dict[str, Any]:
    """Perturb ``hparams`` in place (and return it)."""
    rng = rng or random
    table = get_hp_range_definition()
    for key in list(hparams.keys()):
        value = hparams[key]
        if isinstance(value, bool):
            continue
        if key != "opt_case" and key not in table:
            continue        # a key without a range entry has nothing to clip against: leave it
        if isinstance(value, float):
            hparams[key] = perturb_float(value, table[key][0], table[key][-1], rng, factors)
        elif isinstance(value, int):
            if key == "batch_size":
                hparams[key] = perturb_int(value, BATCH_SIZE_OFFSET,
                                           table[key][-1] + BATCH_SIZE_OFFSET, rng, factors)
            else:
                hparams[key] = perturb_int(value, table[key][0], table[key][-1], rng, factors)
        elif key == "opt_case":
            name = value["optimizer"]
            lr_list = table["lr"][name]
            value["lr"] = perturb_float(value["lr"], lr_list[0], lr_list[-1], rng, factors)
            if name in ("Momentum", "RMSProp"):
                value["momentum"] = perturb_float(value["momentum"], table["momentum"][0],
                                                  table["momentum"][-1], rng, factors)
            if name == "RMSProp":
                value["grad_decay"] = perturb_float(value["grad_decay"], table["grad_decay"][0],
                                                    table["grad_decay"][-1], rng, factors)
        elif isinstance(value, str):
            if key not in FROZEN_STRING_KEYS and key in table:
                hparams[key] = rng.choice(table[key])
    return hparams
