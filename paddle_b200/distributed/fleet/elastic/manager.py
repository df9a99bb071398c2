"""ElasticManager on a file-based membership store (reference: fleet/elastic/manager.py).

Nodes register ``nodes/<host>`` with a heartbeat timestamp (TTL 10 s, refreshed every TTL/2), watch the member set,
and classify changes into HOLD (fewer than min nodes), RESTART (membership changed within [min, max]), ERROR,
COMPLETED, EXIT - the same state machine as the reference's etcd implementation; training-state continuity is the
user's job (reload the last checkpoint), as in the reference."""
from __future__ import annotations

import json
import os
import threading
import time

ELASTIC_TTL = 10
ELASTIC_EXIT_CODE = 101
ELASTIC_AUTO_PARALLEL_EXIT_CODE = 102


class ElasticStatus:
    COMPLETED = "completed"
    ERROR = "error"
    HOLD = "hold"
    RESTART = "restart"
    EXIT = "exit"


class ElasticLevel:
    FAULT_TOLERANCE = 1
    ELASTIC = 2


class LauncherInterface:
    def __init__(self, args):
        self.args = args
        self.procs = []

    def launch(self):
        raise NotImplementedError

    def stop(self):
        raise NotImplementedError

    def watch(self):
        raise NotImplementedError


class FileStore:
    """Tiny KV with TTL on a shared directory (stands in for etcd in tests and single-host jobs)."""

    def __init__(self, root):
        self.root = root
        os.makedirs(root, exist_ok=True)

    def _p(self, key):
        return os.path.join(self.root, key.strip("/").replace("/", "__"))

    def put(self, key, value, ttl=None):
        tmp = self._p(key) + ".tmp"
        with open(tmp, "w") as f:
            json.dump({"v": value, "exp": time.time() + ttl if ttl else None}, f)
        os.replace(tmp, self._p(key))

    def get(self, key):
        try:
            with open(self._p(key)) as f:
                d = json.load(f)
        except (OSError, ValueError):
            return None
        if d["exp"] is not None and d["exp"] < time.time():
            return None
        return d["v"]

    def get_prefix(self, prefix):
        pre = prefix.strip("/").replace("/", "__")
        out = {}
        for fn in os.listdir(self.root):
            if fn.startswith(pre) and not fn.endswith(".tmp"):
                v = self.get(fn.replace("__", "/"))
                if v is not None:
                    out[fn.replace("__", "/")] = v
        return out

    def delete(self, key):
        try:
            os.remove(self._p(key))
        except OSError:
            pass


class ElasticManager:
    def __init__(self, args, store=None):
        self.args = args
        np_spec = str(getattr(args, "np", None) or getattr(args, "nnodes", "1"))
        if ":" in np_spec:
            lo, hi = np_spec.split(":")
            self.min_np, self.max_np = int(lo), int(hi)
        else:
            self.min_np = self.max_np = int(np_spec)
        self.elastic_level = ElasticLevel.ELASTIC if self.min_np != self.max_np else ElasticLevel.FAULT_TOLERANCE
        self.host = getattr(args, "host", None) or os.environ.get("POD_IP", "127.0.0.1")
        self.job_id = getattr(args, "job_id", "default")
        self.elastic_timeout = int(getattr(args, "elastic_timeout", 30))
        self.store = store or FileStore(os.environ.get("PADDLE_ELASTIC_STORE", f"/tmp/paddle_b200_elastic_{self.job_id}"))
        self.prefix = f"/paddle/{self.job_id}"
        self.node_key = f"{self.prefix}/nodes/{self.host}"
        self.enable = self.max_np > 1 or self.elastic_level == ElasticLevel.ELASTIC
        self.stopped = False
        self.hosts = []
        self.launcher = None
        self._hb = None
        self.register()

    def register(self):
        self.store.put(self.node_key, self.host, ttl=ELASTIC_TTL)

        def beat():
            while not self.stopped:
                self.store.put(self.node_key, self.host, ttl=ELASTIC_TTL)
                time.sleep(ELASTIC_TTL / 2)
        self._hb = threading.Thread(target=beat, daemon=True)
        self._hb.start()

    def _current_hosts(self):
        return sorted(self.store.get_prefix(f"{self.prefix}/nodes").values())

    def _match(self, hosts=None):
        hosts = hosts if hosts is not None else self._current_hosts()
        if self.elastic_level == ElasticLevel.FAULT_TOLERANCE:
            return len(hosts) == self.max_np
        return self.min_np <= len(hosts) <= self.max_np

    def wait(self):
        """Block until the member set is acceptable; returns False on timeout."""
        t0 = time.time()
        while not self.stopped:
            hosts = self._current_hosts()
            if self._match(hosts):
                self.hosts = hosts
                return True
            if time.time() - t0 > self.elastic_timeout and len(hosts) >= self.min_np:
                self.hosts = hosts
                return True
            if time.time() - t0 > 10 * self.elastic_timeout:
                return False
            time.sleep(1)
        return False

    def endpoints(self, start_port=6170):
        return [f"{h}:{start_port}" for h in self.hosts]

    def run(self, launcher):
        self.launcher = launcher
        self.launcher.launch()

    def watch(self):
        """One poll of the state machine."""
        if self.stopped:
            return ElasticStatus.EXIT
        if self.launcher is not None:
            ret = self.launcher.watch()
            if ret is not None:
                if ret == 0:
                    return ElasticStatus.COMPLETED
                return ElasticStatus.RESTART if self.elastic_level >= ElasticLevel.FAULT_TOLERANCE else ElasticStatus.ERROR
        hosts = self._current_hosts()
        if hosts != self.hosts:
            if len(hosts) < self.min_np:
                return ElasticStatus.HOLD
            return ElasticStatus.RESTART
        return None

    def exit(self, completed=False):
        self.stopped = True
        if self.launcher is not None:
            self.launcher.stop()
        self.store.delete(self.node_key)
        if completed:
            self.store.put(f"{self.prefix}/completed", "1")
