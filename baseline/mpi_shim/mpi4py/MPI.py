"""The subset of ``mpi4py.MPI`` used by the reference drivers, over local TCP sockets.

Calls covered: ``COMM_WORLD.Get_rank/Get_size/Isend/Irecv/bcast/Barrier``,
``Request.Wait/test/Cancel`` and the static ``Request.Waitany/Waitall``, ``Status.Get_source/
Get_tag``, ``DOUBLE``, ``Finalize``.  Ranks are processes started by ``baseline/mpirun_shim.py``
(environment EH_MPI_RANK / EH_MPI_SIZE / EH_MPI_PORT0).  Sends are eager/buffered; receives
match on (source, tag) in posting order like MPI.  This exists only to time the reference's own
CPU code path when no MPI runtime can be installed; it is not part of the product.
"""
import os
import pickle
import socket
import struct
import threading
import time

import numpy as np

DOUBLE = "d"
_RANK = int(os.environ.get("EH_MPI_RANK", "0"))
_SIZE = int(os.environ.get("EH_MPI_SIZE", "1"))
_PORT0 = int(os.environ.get("EH_MPI_PORT0", "47000"))
_CTRL = -12345


class Status(object):
    def __init__(self):
        self.source, self.tag = -1, -1

    def Get_source(self):
        return self.source

    def Get_tag(self):
        return self.tag


class _Mailbox(object):
    def __init__(self):
        self.lock = threading.Condition()
        self.msgs = []          # (src, tag, payload)

    def put(self, src, tag, payload):
        with self.lock:
            self.msgs.append((src, tag, payload))
            self.lock.notify_all()

    def take(self, src, tag, block):
        with self.lock:
            while True:
                for i, (s, t, p) in enumerate(self.msgs):
                    if s == src and t == tag:
                        del self.msgs[i]
                        return p
                if not block:
                    return None
                self.lock.wait(0.05)


_box = _Mailbox()
_out = {}
_out_lock = threading.Lock()


def _recv_exact(conn, n):
    buf = bytearray()
    while len(buf) < n:
        chunk = conn.recv(n - len(buf))
        if not chunk:
            return None
        buf.extend(chunk)
    return bytes(buf)


def _serve(conn):
    while True:
        hdr = _recv_exact(conn, 16)
        if hdr is None:
            return
        src, tag, n = struct.unpack("<iiq", hdr)
        payload = _recv_exact(conn, n) if n else b""
        _box.put(src, tag, payload)


def _listen():
    srv = socket.socket()
    srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    srv.bind(("127.0.0.1", _PORT0 + _RANK))
    srv.listen(_SIZE + 4)
    while True:
        conn, _ = srv.accept()
        conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        threading.Thread(target=_serve, args=(conn,), daemon=True).start()


if _SIZE > 1:
    threading.Thread(target=_listen, daemon=True).start()


# EH_MPI_LINK_LATENCY_MS=L: a message from rank r to rank 0 spends r * L milliseconds on its link (tests use it to
# give the master a deterministic arrival order where the reference has no delay model of its own).
_LINK_MS = float(os.environ.get("EH_MPI_LINK_LATENCY_MS", "0") or 0)


def _send(dest, tag, payload):
    if _LINK_MS > 0 and dest == 0 and _RANK != 0:
        time.sleep(_RANK * _LINK_MS * 1e-3)
    with _out_lock:
        s = _out.get(dest)
        if s is None:
            for _ in range(600):
                try:
                    s = socket.create_connection(("127.0.0.1", _PORT0 + dest))
                    break
                except OSError:
                    time.sleep(0.05)
            if s is None:
                raise RuntimeError("mpi shim: rank %d cannot reach rank %d on port %d" % (_RANK, dest, _PORT0 + dest))
            s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            _out[dest] = s
        s.sendall(struct.pack("<iiq", _RANK, tag, len(payload)) + payload)


class Request(object):
    def __init__(self, kind=None, buf=None, source=None, tag=None):
        self.kind, self.buf, self.source, self.tag = kind, buf, source, tag
        self.done = kind in (None, "send")

    def _poll(self, block=False):
        if self.done:
            return True
        p = _box.take(self.source, self.tag, block)
        if p is None:
            return False
        arr = np.frombuffer(p, dtype=np.float64)
        self.buf[...] = arr.reshape(self.buf.shape)
        self.done = True
        return True

    def Wait(self, status=None):
        self._poll(True)
        if status is not None:
            status.source, status.tag = self.source, self.tag
        return True

    def test(self):
        return (self._poll(False), None)

    def Cancel(self):
        self.done = True

    @staticmethod
    def Waitany(requests, status=None):
        while True:
            for i, r in enumerate(requests):
                if r.kind == "recv" and not r.done and r._poll(False):
                    if status is not None:
                        status.source, status.tag = r.source, r.tag
                    return i
            if not any(r.kind == "recv" and not r.done for r in requests):
                return -1
            time.sleep(0.00005)

    @staticmethod
    def Waitall(requests, statuses=None):
        for r in requests:
            r.Wait()
        return True


class _Comm(object):
    def Get_rank(self):
        return _RANK

    def Get_size(self):
        return _SIZE

    def Isend(self, spec, dest, tag=0):
        buf = spec[0] if isinstance(spec, (list, tuple)) else spec
        _send(dest, tag, np.ascontiguousarray(buf, dtype=np.float64).tobytes())
        return Request("send")

    def Irecv(self, spec, source, tag=0):
        buf = spec[0] if isinstance(spec, (list, tuple)) else spec
        return Request("recv", buf, source, tag)

    def bcast(self, obj, root=0):
        if _SIZE == 1:
            return obj
        if _RANK == root:
            data = pickle.dumps(obj)
            for r in range(_SIZE):
                if r != root:
                    _send(r, _CTRL, data)
            return obj
        return pickle.loads(_box.take(root, _CTRL, True))

    _bar = 0

    def Barrier(self):
        if _SIZE == 1:
            return
        _Comm._bar += 1
        tag = _CTRL - _Comm._bar
        if _RANK == 0:
            for r in range(1, _SIZE):
                _box.take(r, tag, True)
            for r in range(1, _SIZE):
                _send(r, tag, b"")
        else:
            _send(0, tag, b"")
            _box.take(0, tag, True)


COMM_WORLD = _Comm()


def Finalize():
    time.sleep(0.05)
