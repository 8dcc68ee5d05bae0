"""Rendezvous stores used to hand the NCCL unique id from rank 0 to its peers.

ref: key-value ``Store{set,get,add,wait}`` paddle/phi/core/distributed/store/store.h:23-38, socket
``TCPStore(host, port, is_master, num_workers, timeout)`` store/tcp_store.h:120-146; the id is
published under ``ProcessGroupNCCL/nccl_ids/<gid>/0`` (process_group_nccl.cc:559-571).

``TCPStore`` is a small pure-Python equivalent (rank 0 listens).  ``FileStore`` covers single-node
launchers whose master port is already taken by the launcher itself (torchrun's agent store).
"""
from __future__ import annotations

import os
import socket
import struct
import threading
import time


class FileStore:
    def __init__(self, path: str, timeout: float = 900.0):
        self.path = path
        self.timeout = timeout
        os.makedirs(path, exist_ok=True)

    def _file(self, key: str) -> str:
        return os.path.join(self.path, key.replace("/", "__"))

    def set(self, key: str, value: bytes):
        tmp = self._file(key) + ".tmp.%d" % os.getpid()
        with open(tmp, "wb") as f:
            f.write(value)
        os.replace(tmp, self._file(key))  # atomic publish

    def get(self, key: str) -> bytes:
        self.wait(key)
        with open(self._file(key), "rb") as f:
            return f.read()

    def wait(self, key: str):
        t0 = time.time()
        while not os.path.exists(self._file(key)):
            if time.time() - t0 > self.timeout:
                raise TimeoutError("FileStore: key %r not published within %.0f s" % (key, self.timeout))
            time.sleep(0.005)

    def add(self, key: str, amount: int) -> int:
        # single-node helper: serialise through an exclusive lock file
        lock = self._file(key) + ".lock"
        while True:
            try:
                fd = os.open(lock, os.O_CREAT | os.O_EXCL | os.O_WRONLY)
                break
            except FileExistsError:
                time.sleep(0.002)
        try:
            cur = int(open(self._file(key)).read()) if os.path.exists(self._file(key)) else 0
            cur += amount
            self.set(key, str(cur).encode())
            return cur
        finally:
            os.close(fd)
            os.unlink(lock)


class TCPStore:
    """Minimal TCP key-value store: the master rank serves, every rank (master included) is a client."""

    _SET, _GET, _ADD = 0, 1, 2

    def __init__(self, host: str, port: int, is_master: bool, world_size: int = 1, timeout: float = 900.0):
        self.host, self.port, self.timeout = host, int(port), timeout
        self._server = None
        if is_master:
            self._data: dict = {}
            self._cv = threading.Condition()
            self._server = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
            self._server.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            self._server.bind((host, self.port))
            self._server.listen(max(world_size, 8))
            threading.Thread(target=self._serve, daemon=True).start()
        self._sock = self._connect()

    # ---- server
    def _serve(self):
        while True:
            try:
                conn, _ = self._server.accept()
            except OSError:
                return
            threading.Thread(target=self._handle, args=(conn,), daemon=True).start()

    def _handle(self, conn):
        try:
            while True:
                hdr = self._recv_exact(conn, 9)
                if hdr is None:
                    return
                op, klen, vlen = struct.unpack("!BII", hdr)
                key = self._recv_exact(conn, klen).decode()
                val = self._recv_exact(conn, vlen) if vlen else b""
                if op == self._SET:
                    with self._cv:
                        self._data[key] = val
                        self._cv.notify_all()
                    conn.sendall(struct.pack("!I", 0))
                elif op == self._GET:
                    with self._cv:
                        ok = self._cv.wait_for(lambda: key in self._data, timeout=self.timeout)
                        out = self._data[key] if ok else b""
                    conn.sendall(struct.pack("!I", len(out)) + out)
                elif op == self._ADD:
                    with self._cv:
                        cur = int(self._data.get(key, b"0")) + struct.unpack("!q", val)[0]
                        self._data[key] = str(cur).encode()
                        self._cv.notify_all()
                    out = str(cur).encode()
                    conn.sendall(struct.pack("!I", len(out)) + out)
        except (ConnectionError, OSError):
            return
        finally:
            conn.close()

    @staticmethod
    def _recv_exact(conn, n):
        buf = b""
        while len(buf) < n:
            chunk = conn.recv(n - len(buf))
            if not chunk:
                return None
            buf += chunk
        return buf

    # ---- client
    def _connect(self):
        t0 = time.time()
        while True:
            try:
                s = socket.create_connection((self.host, self.port), timeout=5.0)
                s.settimeout(self.timeout)
                return s
            except OSError:
                if time.time() - t0 > self.timeout:
                    raise
                time.sleep(0.05)

    def _request(self, op, key: str, val: bytes = b"") -> bytes:
        k = key.encode()
        self._sock.sendall(struct.pack("!BII", op, len(k), len(val)) + k + val)
        n = struct.unpack("!I", self._recv_exact(self._sock, 4))[0]
        return self._recv_exact(self._sock, n) if n else b""

    def set(self, key: str, value: bytes):
        self._request(self._SET, key, value)

    def get(self, key: str) -> bytes:
        return self._request(self._GET, key)

    def wait(self, key: str):
        self.get(key)

    def add(self, key: str, amount: int) -> int:
        return int(self._request(self._ADD, key, struct.pack("!q", amount)))

    def close(self):
        try:
            self._sock.close()
            if self._server:
                self._server.close()
        except OSError:
            pass
