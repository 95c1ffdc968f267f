"""The process-wide gate between CUDA-graph captures and steps (utils/capture_gate.py): steps overlap each other, a
capture waits for running steps and runs alone, steps that arrive during a capture wait for it."""
import threading
import time

from distributedtf_b200.utils.capture_gate import StepGate


def test_steps_overlap_and_capture_runs_alone():
    gate = StepGate()
    log, lock = [], threading.Lock()
    inside = {"steps": 0, "capture": 0, "max_steps": 0, "overlap": False}

    def step(i):
        gate.enter_step()
        with lock:
            inside["steps"] += 1
            inside["max_steps"] = max(inside["max_steps"], inside["steps"])
            inside["overlap"] |= inside["capture"] > 0
        time.sleep(0.02)
        with lock:
            inside["steps"] -= 1
        gate.exit_step()

    def capture():
        gate.enter_capture()
        with lock:
            inside["capture"] += 1
            inside["overlap"] |= inside["steps"] > 0
        time.sleep(0.03)
        with lock:
            inside["capture"] -= 1
        gate.exit_capture()
        log.append("captured")

    threads = [threading.Thread(target=step, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    time.sleep(0.005)                         # the steps are in flight when the capture asks for the gate
    threads.append(threading.Thread(target=capture))
    threads[-1].start()
    time.sleep(0.002)
    late = [threading.Thread(target=step, args=(10 + i,)) for i in range(3)]   # arrive while the capture waits / runs
    for t in late:
        t.start()
    for t in threads + late:
        t.join(5)
        assert not t.is_alive()
    assert inside["max_steps"] >= 2           # steps ran side by side
    assert not inside["overlap"]              # never a step and a capture at once
    assert log == ["captured"]


def test_two_captures_take_turns():
    gate = StepGate()
    active, worst = [0], [0]

    def capture():
        gate.enter_capture()
        active[0] += 1
        worst[0] = max(worst[0], active[0])
        time.sleep(0.01)
        active[0] -= 1
        gate.exit_capture()

    ts = [threading.Thread(target=capture) for _ in range(5)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(5)
    assert worst[0] == 1
