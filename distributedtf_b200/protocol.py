"""Wire protocol between the PBT master and its workers.

Parity: reference ``constants.py:5-12`` (enum values) and SURVEY §2.6 (tuple
shapes).  The integer values are part of the protocol and must not change; the
messages are plain tuples whose first element is the instruction, so they can be
pickled by any of the control-plane transports in ``parallel/``.

Messages (master → worker unless noted)::

    (ADD_GRAPHS, [hparams, ...], first_member_id, explore_only)
    (TRAIN, epochs_this_round, epochs_per_round * rounds)
    (GET,)                       -> worker replies [[id, accuracy, hparams], ...]
    (SET, [[id, accuracy, hparams(, extras)], ...])
    (EXPLORE,)
    (GET_PROFILING_INFO,)        -> worker replies [train_seconds, explore_seconds]
    (EXIT,)
"""
from __future__ import annotations

import enum
from typing import Any, Sequence, Tuple


class WorkerInstruction(enum.Enum):
    ADD_GRAPHS = 0
    EXIT = 1
    TRAIN = 2
    GET = 3
    SET = 4
    EXPLORE = 5
    GET_PROFILING_INFO = 6


# Instructions that make the worker send something back to the master.
REPLYING = frozenset({WorkerInstruction.GET, WorkerInstruction.GET_PROFILING_INFO})


def add_graphs(hparam_list: Sequence[dict], first_id: int, explore_only: bool) -> Tuple:
    return (WorkerInstruction.ADD_GRAPHS, list(hparam_list), int(first_id), bool(explore_only))


def train(epochs: int, total_epochs: int) -> Tuple:
    return (WorkerInstruction.TRAIN, int(epochs), int(total_epochs))


def get() -> Tuple:
    return (WorkerInstruction.GET,)


def set_values(values: Sequence[Sequence[Any]]) -> Tuple:
    return (WorkerInstruction.SET, [list(v) for v in values])


def explore() -> Tuple:
    return (WorkerInstruction.EXPLORE,)


def get_profiling_info() -> Tuple:
    return (WorkerInstruction.GET_PROFILING_INFO,)


def exit_() -> Tuple:
    return (WorkerInstruction.EXIT,)


def validate(msg: Any) -> WorkerInstruction:
    """Check the shape of a decoded message; returns its instruction."""
    if not isinstance(msg, tuple) or not msg or not isinstance(msg[0], WorkerInstruction):
        raise ValueError(f"malformed control message: {msg!r}")
    inst = msg[0]
    want = {
        WorkerInstruction.ADD_GRAPHS: 4,
        WorkerInstruction.TRAIN: 3,
        WorkerInstruction.GET: 1,
        WorkerInstruction.SET: 2,
        WorkerInstruction.EXPLORE: 1,
        WorkerInstruction.GET_PROFILING_INFO: 1,
        WorkerInstruction.EXIT: 1,
    }[inst]
    if len(msg) != want:
        raise ValueError(f"{inst.name} expects a {want}-tuple, got {len(msg)}")
    return inst
