"""One gate per process between CUDA-graph captures and everything else that talks to the CUDA runtime.

Steps of different members (threads of one worker) run concurrently; a CUDA-graph capture runs alone.  The master of
an in-process run (``transport="local"``: master and workers are threads of one process) polls the device score
board with small device-to-host copies; issued while a worker thread was capturing, those calls failed with
``cudaErrorStreamCaptureUnsupported`` and invalidated the capture (seen at population 50 of the reference's sweep
protocol, where hundreds of captures meet a 10 kHz poll).  So the board snapshot enters the gate like a step does.
"""
from __future__ import annotations

import threading


class StepGate:
    def __init__(self):
        self._cv = threading.Condition()
        self._active = 0
        self._capturing = False

    def enter_step(self) -> None:
        with self._cv:
            while self._capturing:
                self._cv.wait()
            self._active += 1

    def exit_step(self) -> None:
        with self._cv:
            self._active -= 1
            if self._active == 0:
                self._cv.notify_all()

    def enter_capture(self) -> None:
        with self._cv:
            while self._capturing:
                self._cv.wait()
            self._capturing = True
            while self._active > 0:
                self._cv.wait()

    def exit_capture(self) -> None:
        with self._cv:
            self._capturing = False
            self._cv.notify_all()


GATE = StepGate()
