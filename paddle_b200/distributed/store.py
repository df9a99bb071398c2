"""TCPStore on the native runtime (reference: paddle/phi/core/distributed/store/tcp_store.h:40-151, exposed to Python
as core.TCPStore / create_or_get_global_tcp_store).  Used for rendezvous of things torch.distributed does not
carry: symmetric-memory handle exchange bookkeeping, elastic membership, user key/values."""
from __future__ import annotations

import ctypes
import os

from .. import lib


class TCPStore:
    def __init__(self, host, port, is_master=False, world_size=1, timeout=900):
        rt = lib.runtime.require()
        self._rt = rt
        self._h = rt.pdb200_store_create(str(host).encode(), int(port), int(bool(is_master)), int(world_size),
                                         int(timeout * 1000))
        if not self._h:
            raise RuntimeError(f"TCPStore: {rt.pdb200_rt_last_error().decode()}")
        self.host, self.world_size = host, world_size
        self.port = rt.pdb200_store_port(self._h)

    def set(self, key, value):
        data = value if isinstance(value, (bytes, bytearray)) else str(value).encode()
        if self._rt.pdb200_store_set(self._h, key.encode(), data, len(data)) != 0:
            raise RuntimeError("TCPStore.set failed")

    def get(self, key, timeout=None):
        cap = 1 << 16
        tmo = -1 if timeout is None else int(timeout * 1000)
        while True:
            buf = ctypes.create_string_buffer(cap)
            n = self._rt.pdb200_store_get(self._h, key.encode(), buf, cap, tmo)
            if n >= 0:
                return buf.raw[:n]
            if n == -2:
                raise TimeoutError(f"TCPStore.get({key!r}) timed out")
            if n < -16:
                cap = -(n + 16)
                continue
            raise RuntimeError("TCPStore.get failed")

    def add(self, key, value=1):
        v = self._rt.pdb200_store_add(self._h, key.encode(), int(value))
        if v == -(2 ** 63):
            raise RuntimeError("TCPStore.add failed")
        return v

    def check(self, key):
        return self._rt.pdb200_store_check(self._h, key.encode()) == 1

    def wait(self, key, timeout=None):
        rc = self._rt.pdb200_store_wait(self._h, key.encode(), -1 if timeout is None else int(timeout * 1000))
        if rc == -2:
            raise TimeoutError(f"TCPStore.wait({key!r}) timed out")
        if rc != 0:
            raise RuntimeError("TCPStore.wait failed")

    def delete_key(self, key):
        return self._rt.pdb200_store_delete(self._h, key.encode()) == 1

    def close(self):
        if getattr(self, "_h", None):
            self._rt.pdb200_store_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


_global_store = None


def create_or_get_global_tcp_store():
    """Store on ``PADDLE_MASTER`` (port + 1 so it does not collide with the process-group rendezvous store)."""
    global _global_store
    if _global_store is None:
        from .parallel import ParallelEnv, _master_endpoint
        env = ParallelEnv()
        host, port = _master_endpoint(env)
        _global_store = TCPStore(host, port + 1, is_master=(env.rank == 0), world_size=env.world_size)
    return _global_store
