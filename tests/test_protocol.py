"""Wire protocol: enum values and tuple shapes (reference constants.py:5-12)."""
import pickle

import pytest

from distributedtf_b200 import protocol
from distributedtf_b200.protocol import WorkerInstruction as WI


def test_enum_values_are_the_reference_protocol():
    assert [(m.name, m.value) for m in WI] == [
        ("ADD_GRAPHS", 0), ("EXIT", 1), ("TRAIN", 2), ("GET", 3), ("SET", 4),
        ("EXPLORE", 5), ("GET_PROFILING_INFO", 6)]


def test_message_shapes():
    assert protocol.add_graphs([{"a": 1}], 3, True) == (WI.ADD_GRAPHS, [{"a": 1}], 3, True)
    assert protocol.train(1, 20) == (WI.TRAIN, 1, 20)
    assert protocol.get() == (WI.GET,)
    assert protocol.set_values([[0, 0.5, {}]]) == (WI.SET, [[0, 0.5, {}]])
    assert protocol.explore() == (WI.EXPLORE,)
    assert protocol.get_profiling_info() == (WI.GET_PROFILING_INFO,)
    assert protocol.exit_() == (WI.EXIT,)


@pytest.mark.parametrize("msg", [protocol.add_graphs([], 0, False), protocol.train(2, 4),
                                 protocol.get(), protocol.set_values([]), protocol.explore(),
                                 protocol.get_profiling_info(), protocol.exit_()])
def test_pickle_round_trip_and_validate(msg):
    back = pickle.loads(pickle.dumps(msg))
    assert back == msg
    assert protocol.validate(back) is msg[0]


@pytest.mark.parametrize("bad", [None, (), ("TRAIN", 1, 2), (WI.TRAIN, 1), (WI.GET, 1), [WI.GET]])
def test_validate_rejects_malformed(bad):
    with pytest.raises(ValueError):
        protocol.validate(bad)
