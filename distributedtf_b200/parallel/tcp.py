"""Star-topology TCP control plane: the master listens, every worker connects.

Replaces Open MPI + mpi4py (reference ``README.md:7-11``, ``main_manager.py:46``)
on an image that has neither ``mpirun`` nor mpi4py.  Only master<->worker
traffic exists in the protocol, so a star is sufficient.  Frames
are ``8-byte big-endian length + pickle``.  Per-connection ordering is TCP's, so
the FIFO guarantee of ``Comm`` holds.  The master keeps one reader thread per
worker that parks decoded frames on a per-source queue, so ``recv(source=i)``
never sees another worker's reply.
"""
from __future__ import annotations

import queue
import socket
import struct
import threading
import time
from typing import Any, Dict, Optional

from .comm_base import Comm, CommTimeout, Request, dumps, loads

_HDR = struct.Struct("!Q")


def _send_frame(sock: socket.socket, payload: bytes) -> None:
    sock.sendall(_HDR.pack(len(payload)) + payload)


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    chunks = []
    while n:
        b = sock.recv(min(n, 1 << 20))
        if not b:
            raise ConnectionError("control-plane peer closed the connection")
        chunks.append(b)
        n -= len(b)
    return b"".join(chunks)


def _recv_frame(sock: socket.socket) -> bytes:
    (n,) = _HDR.unpack(_recv_exact(sock, _HDR.size))
    return _recv_exact(sock, n)


class TcpMasterComm(Comm):
    """Endpoint 0.  ``size`` counts the master, as in ``MPI.COMM_WORLD``."""

    def __init__(self, host: str, port: int, size: int, accept_timeout: float = 300.0):
        self._size = size
        self._srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self._srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self._srv.bind((host, port))
        self._srv.listen(size)
        self.port = self._srv.getsockname()[1]
        self._socks: Dict[int, socket.socket] = {}
        self._inbox: Dict[int, "queue.Queue[Any]"] = {r: queue.Queue() for r in range(1, size)}
        self._send_locks: Dict[int, threading.Lock] = {r: threading.Lock() for r in range(1, size)}
        self._accept_timeout = accept_timeout
        self._accepted = threading.Event()
        self._threads = []
        t = threading.Thread(target=self._accept_all, name="dtf-tcp-accept", daemon=True)
        t.start()
        self._threads.append(t)

    def _accept_all(self) -> None:
        self._srv.settimeout(self._accept_timeout)
        while len(self._socks) < self._size - 1:
            conn, _ = self._srv.accept()
            conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            rank = loads(_recv_frame(conn))
            self._socks[rank] = conn
            t = threading.Thread(target=self._reader, args=(rank, conn),
                                 name=f"dtf-tcp-rx{rank}", daemon=True)
            t.start()
            self._threads.append(t)
        self._accepted.set()

    def _reader(self, rank: int, conn: socket.socket) -> None:
        try:
            while True:
                self._inbox[rank].put(_recv_frame(conn))
        except (ConnectionError, OSError):
            self._inbox[rank].put(None)  # poison: peer gone

    def wait_for_workers(self) -> None:
        if not self._accepted.wait(self._accept_timeout):
            raise CommTimeout("not all workers connected to the master")

    def Get_rank(self) -> int:
        return 0

    def Get_size(self) -> int:
        return self._size

    def isend(self, obj: Any, dest: int) -> Request:
        self.wait_for_workers()
        payload = dumps(obj)
        with self._send_locks[dest]:
            _send_frame(self._socks[dest], payload)
        return Request()

    def recv(self, source: int, timeout: Optional[float] = None) -> Any:
        try:
            frame = self._inbox[source].get(timeout=timeout)
        except queue.Empty:
            raise CommTimeout(f"master: no reply from worker {source} within {timeout}s") from None
        if frame is None:
            raise ConnectionError(f"worker {source} disconnected")
        return loads(frame)

    def close(self) -> None:
        for s in self._socks.values():
            try:
                s.close()
            except OSError:
                pass
        self._srv.close()


class TcpWorkerComm(Comm):
    def __init__(self, host: str, port: int, rank: int, size: int, connect_timeout: float = 300.0):
        if rank < 1:
            raise ValueError("worker endpoints are ranks 1..size-1")
        self._rank, self._size = rank, size
        deadline = time.time() + connect_timeout
        last: Optional[Exception] = None
        while True:
            try:
                self._sock = socket.create_connection((host, port), timeout=10.0)
                break
            except OSError as e:  # master not listening yet
                last = e
                if time.time() > deadline:
                    raise CommTimeout(f"worker {rank}: cannot reach master at {host}:{port}: {last}")
                time.sleep(0.05)
        self._sock.settimeout(None)
        self._sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self._lock = threading.Lock()
        _send_frame(self._sock, dumps(rank))

    def Get_rank(self) -> int:
        return self._rank

    def Get_size(self) -> int:
        return self._size

    def isend(self, obj: Any, dest: int) -> Request:
        if dest != 0:
            raise ValueError("workers only talk to the master (star topology)")
        with self._lock:
            _send_frame(self._sock, dumps(obj))
        return Request()

    def recv(self, source: int = 0, timeout: Optional[float] = None) -> Any:
        if source != 0:
            raise ValueError("workers only hear from the master (star topology)")
        self._sock.settimeout(timeout)
        try:
            return loads(_recv_frame(self._sock))
        except socket.timeout:
            raise CommTimeout(f"worker {self._rank}: master silent for {timeout}s") from None
        finally:
            self._sock.settimeout(None)

    def close(self) -> None:
        try:
            self._sock.close()
        except OSError:
            pass
