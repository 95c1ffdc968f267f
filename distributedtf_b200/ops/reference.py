"""fp32 reference restatements: TF-1.x optimizer rules, regularisers, LR schedule,
initialisers.  These are the oracles the CUDA kernels are tested against and the
update rules of the restated (torch-eager) baseline.

Parity: reference ``resnet_run_loop.py:552-586`` / ``mnist_model.py:27-60`` (six
optimizers with TF-1.x defaults), ``resnet_model.py:95-121`` (initialisers and
regularisers), ``resnet_run_loop.py:135-173`` + ``cifar10_main.py:192-208`` (LR
schedule).  torch.optim is *not* used: its RMSprop/Adagrad/Adadelta/Adam differ
from the TF-1.x rules in eps placement and slot initialisation.
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

OPT_IDS = {"gd": 0, "Momentum": 1, "Adagrad": 2, "Adadelta": 3, "Adam": 4, "RMSProp": 5}
REG_IDS = {"None": 0, "l1_regularizer": 1, "l2_regularizer": 2, "l1_l2_regularizer": 3}
NUM_SLOTS = {"gd": 0, "Momentum": 1, "Adagrad": 1, "Adadelta": 2, "Adam": 2, "RMSProp": 2}

ADADELTA_RHO, ADADELTA_EPS = 0.95, 1e-8
ADAGRAD_INIT = 0.1
ADAM_B1, ADAM_B2, ADAM_EPS = 0.9, 0.999, 1e-8
RMSPROP_EPS = 1e-10


def slot_init_values(optimizer: str) -> Tuple[float, float]:
    """Initial fill of (slot0, slot1) for each TF-1.x optimizer."""
    if optimizer == "Adagrad":
        return (ADAGRAD_INIT, 0.0)
    if optimizer == "RMSProp":
        return (1.0, 0.0)          # rms slot starts at ones, momentum slot at zeros
    return (0.0, 0.0)


def regularizer_grad(w: torch.Tensor, regularizer: str, weight_decay: float) -> torch.Tensor:
    """d/dw of l1: wd*sum|w| ; l2: wd*sum(w^2)/2 (tf.nn.l2_loss) ; l1_l2: both."""
    if regularizer == "l1_regularizer":
        return weight_decay * torch.sign(w)
    if regularizer == "l2_regularizer":
        return weight_decay * w
    if regularizer == "l1_l2_regularizer":
        return weight_decay * torch.sign(w) + weight_decay * w
    return torch.zeros_like(w)


def regularizer_loss(ws: Sequence[torch.Tensor], regularizer: str, weight_decay: float) -> torch.Tensor:
    total = torch.zeros((), dtype=torch.float32, device=ws[0].device if ws else "cpu")
    for w in ws:
        if regularizer in ("l1_regularizer", "l1_l2_regularizer"):
            total = total + weight_decay * w.abs().sum()
        if regularizer in ("l2_regularizer", "l1_l2_regularizer"):
            total = total + weight_decay * 0.5 * (w * w).sum()
    return total


@torch.no_grad()
def apply_update(optimizer: str, w: torch.Tensor, g: torch.Tensor, s0: Optional[torch.Tensor],
                 s1: Optional[torch.Tensor], *, lr: float, step: int, momentum: float = 0.0,
                 grad_decay: float = 0.9) -> None:
    """One TF-1.x ``Apply<Optimizer>`` in place.  ``step`` is the 1-based update count
    (Adam's bias correction uses beta^step)."""
    if optimizer == "gd":
        w.sub_(lr * g)
    elif optimizer == "Momentum":
        s0.mul_(momentum).add_(g)
        w.sub_(lr * s0)
    elif optimizer == "Adagrad":
        s0.add_(g * g)
        w.sub_(lr * g / torch.sqrt(s0))
    elif optimizer == "Adadelta":
        s0.mul_(ADADELTA_RHO).add_((1 - ADADELTA_RHO) * g * g)
        upd = torch.sqrt(s1 + ADADELTA_EPS) / torch.sqrt(s0 + ADADELTA_EPS) * g
        s1.mul_(ADADELTA_RHO).add_((1 - ADADELTA_RHO) * upd * upd)
        w.sub_(lr * upd)
    elif optimizer == "Adam":
        lr_t = lr * math.sqrt(1 - ADAM_B2 ** step) / (1 - ADAM_B1 ** step)
        s0.mul_(ADAM_B1).add_((1 - ADAM_B1) * g)
        s1.mul_(ADAM_B2).add_((1 - ADAM_B2) * g * g)
        w.sub_(lr_t * s0 / (torch.sqrt(s1) + ADAM_EPS))
    elif optimizer == "RMSProp":
        s0.mul_(grad_decay).add_((1 - grad_decay) * g * g)
        s1.mul_(momentum).add_(lr * g / torch.sqrt(s0 + RMSPROP_EPS))
        w.sub_(s1)
    else:
        raise RuntimeError("Hyper-parameter optimizer is wrong!")


# --------------------------------------------------------------------------- LR
def cifar_lr_schedule(hp: Dict[str, Any], num_images: int = 50000, batch_denom: int = 128
                      ) -> Tuple[List[int], List[float]]:
    """(boundaries in steps, values) of the piecewise-constant CIFAR schedule.

    base = lr * batch/128; with decay_steps in {0, 100} the schedule is constant;
    otherwise boundaries sit every ``250*decay_steps/100`` epochs and the value is
    multiplied by ``decay_rate`` at each (ceil(100/decay_steps) - 1 of them).
    """
    bs = int(hp["batch_size"])
    base = float(hp["opt_case"]["lr"]) * bs / batch_denom
    ds = int(hp.get("decay_steps", 0))
    per_epoch = num_images / bs
    if ds == 0 or ds == 100:
        return [int(per_epoch * 250)], [base, base]
    n = int(math.ceil(100 / ds)) - 1
    span = 250 * ds / 100.0
    rates = [1.0]
    bounds = []
    for i in range(n):
        rates.append(float(hp["decay_rate"]) * rates[i])
        bounds.append(int(per_epoch * span * (i + 1)))
    return bounds, [base * r for r in rates]


def piecewise_constant(step: int, boundaries: Sequence[int], values: Sequence[float]) -> float:
    """tf.train.piecewise_constant: values[0] for step <= b0, values[i] for b(i-1) < step <= bi."""
    for b, v in zip(boundaries, values):
        if step <= b:
            return v
    return values[len(boundaries)]


def cifar_lr_at(step: int, hp: Dict[str, Any]) -> float:
    b, v = cifar_lr_schedule(hp)
    return piecewise_constant(step, b, v)


# --------------------------------------------------------------------------- init
def _fans(shape: Sequence[int]) -> Tuple[int, int]:
    """fan_in / fan_out for an OHWI conv kernel [out, kh, kw, in] or a dense [out, in]."""
    if len(shape) == 4:
        rf = shape[1] * shape[2]
        return shape[3] * rf, shape[0] * rf
    return shape[1], shape[0]


def _truncated_normal(shape, std: float, gen: torch.Generator) -> torch.Tensor:
    t = torch.empty(tuple(shape), dtype=torch.float32)
    torch.nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=gen)
    return t


def init_kernel(shape: Sequence[int], initializer: str, gen: torch.Generator) -> torch.Tensor:
    """TF-1.x initialisers for a conv/dense kernel stored [out, (kh, kw,) in]."""
    fan_in, fan_out = _fans(shape)
    if initializer == "glorot_normal":
        return _truncated_normal(shape, math.sqrt(2.0 / (fan_in + fan_out)), gen)
    if initializer == "he_init":
        return _truncated_normal(shape, math.sqrt(2.0 / fan_in), gen)
    if initializer == "orthogonal":
        # tf.orthogonal_initializer flattens to [prod(shape[:-1]), shape[-1]] in its
        # own (HWIO) layout, i.e. rows = kh*kw*in, cols = out.
        rows, cols = fan_in, shape[0]
        a = torch.randn(max(rows, cols), min(rows, cols), generator=gen)
        q, r = torch.linalg.qr(a)
        q = q * torch.sign(torch.diagonal(r))
        if rows < cols:
            q = q.t()
        return q.reshape(rows, cols).t().reshape(tuple(shape)).contiguous()
    limit = math.sqrt(6.0 / (fan_in + fan_out))          # glorot_uniform (layer default)
    return (torch.rand(tuple(shape), generator=gen) * 2 - 1) * limit
