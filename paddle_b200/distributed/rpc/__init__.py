"""paddle.distributed.rpc (reference: python/paddle/distributed/rpc/rpc.py:73-400 over a brpc agent,
paddle/fluid/distributed/rpc/): run a python function on another worker and get its result.

Every worker runs one listener thread; a call is one length-prefixed pickle ``(fn, args, kwargs)`` over a TCP
connection, executed on a small thread pool of the callee, answered with ``(ok, value-or-exception)``.  Workers find
each other through the native TCP store (``paddle_b200.distributed.store.TCPStore``, rank 0 hosts it at
``master_endpoint``): each publishes ``name -> (rank, ip, port)``; ``shutdown`` is a store-counted barrier so that nobody
closes its listener while a peer may still call it."""
from __future__ import annotations

import os
import pickle
import socket
import struct
import threading
from collections import namedtuple
from concurrent.futures import Future, ThreadPoolExecutor

from ..store import TCPStore

WorkerInfo = namedtuple("WorkerInfo", ["name", "rank", "ip", "port"])
_DEFAULT_RPC_TIMEOUT = -1
_state = {"store": None, "self": None, "infos": {}, "server": None, "pool": None, "stop": False, "world": 0}


def _send_msg(sock, payload):
    sock.sendall(struct.pack("!Q", len(payload)) + payload)


def _recv_exact(sock, n):
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(min(n - len(buf), 1 << 20))
        if not chunk:
            raise ConnectionError("rpc peer closed the connection")
        buf += chunk
    return bytes(buf)


def _recv_msg(sock):
    (n,) = struct.unpack("!Q", _recv_exact(sock, 8))
    return _recv_exact(sock, n)


def _serve_one(conn):
    try:
        with conn:
            fn, args, kwargs = pickle.loads(_recv_msg(conn))
            try:
                reply = (True, fn(*args, **kwargs))
            except Exception as e:  # noqa: BLE001  (the exception travels back to the caller)
                reply = (False, e)
            try:
                payload = pickle.dumps(reply)
            except Exception as e:  # noqa: BLE001
                payload = pickle.dumps((False, RuntimeError(f"rpc result is not picklable: {e}")))
            _send_msg(conn, payload)
    except (ConnectionError, OSError):
        pass


def _listen(server):
    while not _state["stop"]:
        try:
            conn, _ = server.accept()
        except OSError:
            break
        _state["pool"].submit(_serve_one, conn)


def _free_endpoint():
    ip = os.environ.get("POD_IP", "127.0.0.1")
    return ip, 0


def init_rpc(name, rank=None, world_size=None, master_endpoint=None):
    """Start this worker's RPC service and learn about the others (reference: rpc.py:73)."""
    rank = int(os.environ["PADDLE_TRAINER_ID"]) if rank is None else int(rank)
    world_size = int(os.environ["PADDLE_TRAINERS_NUM"]) if world_size is None else int(world_size)
    master_endpoint = master_endpoint if master_endpoint is not None else os.environ["PADDLE_MASTER_ENDPOINT"]
    host, port = master_endpoint.split(":")
    ep = os.environ.get("PADDLE_WORKER_ENDPOINT")
    ip, my_port = (ep.split(":")[0], int(ep.split(":")[1])) if ep else _free_endpoint()
    server = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    server.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    server.bind((ip if ip != "127.0.0.1" else "0.0.0.0", my_port))
    server.listen(128)
    my_port = server.getsockname()[1]
    _state.update(stop=False, server=server, pool=ThreadPoolExecutor(max_workers=8, thread_name_prefix="pdb200_rpc"),
                  world=world_size)
    threading.Thread(target=_listen, args=(server,), daemon=True, name="pdb200_rpc_listener").start()
    store = TCPStore(host, int(port), is_master=(rank == 0), world_size=world_size,
                     timeout=int(os.environ.get("FLAGS_stop_check_timeout", "900")))
    me = WorkerInfo(name, rank, ip, my_port)
    store.set(f"rpc/worker/{rank}", pickle.dumps(tuple(me)))
    infos = {}
    for r in range(world_size):
        info = WorkerInfo(*pickle.loads(store.get(f"rpc/worker/{r}")))
        assert info.name not in infos, f"The Worker name must be unique, but name `{info.name}` is repeated."
        infos[info.name] = info
    _state.update(store=store, self=me, infos=infos)
    _barrier("init")


def _barrier(tag):
    store, world = _state["store"], _state["world"]
    n = store.add(f"rpc/barrier/{tag}", 1)
    if n == world:
        store.set(f"rpc/barrier/{tag}/done", b"1")
    store.wait(f"rpc/barrier/{tag}/done")


def _call(to, fn, args, kwargs, timeout):
    info = get_worker_info(to)
    with socket.create_connection((info.ip, info.port), timeout=None if timeout is None or timeout <= 0 else timeout) as s:
        _send_msg(s, pickle.dumps((fn, tuple(args or ()), dict(kwargs or {}))))
        ok, value = pickle.loads(_recv_msg(s))
    if not ok:
        raise value
    return value


def rpc_sync(to, fn, args=None, kwargs=None, timeout=_DEFAULT_RPC_TIMEOUT):
    """Run ``fn(*args, **kwargs)`` on worker ``to`` and return its result (blocking; reference: rpc.py:143).  ``timeout``
    in seconds, <= 0 means none.  An exception raised by ``fn`` is re-raised here."""
    return _call(to, fn, args, kwargs, timeout)


class _RpcFuture(Future):
    def wait(self, timeout=None):
        """The reference's FutureWrapper API: block and return the value."""
        return self.result(timeout)


def rpc_async(to, fn, args=None, kwargs=None, timeout=_DEFAULT_RPC_TIMEOUT):
    """Non-blocking variant: returns a future with ``wait()`` (reference: rpc.py:183)."""
    fut = _RpcFuture()

    def run():
        try:
            fut.set_result(_call(to, fn, args, kwargs, timeout))
        except BaseException as e:  # noqa: BLE001
            fut.set_exception(e)
    threading.Thread(target=run, daemon=True).start()
    return fut


def shutdown():
    """Barrier over all workers, then stop the service (reference: rpc.py:276)."""
    if _state["store"] is None:
        return
    _barrier("shutdown")
    _state["stop"] = True
    try:
        _state["server"].close()
    except OSError:
        pass
    _state["pool"].shutdown(wait=True)
    if _state["self"].rank != 0:
        _state["store"].close()
    else:
        # the master's store must outlive the last waiter of the shutdown barrier
        _barrier_master_linger()
    _state.update(store=None, self=None, infos={}, server=None, pool=None)


def _barrier_master_linger():
    import time
    time.sleep(0.2)
    _state["store"].close()


def get_worker_info(name):
    return _state["infos"][name]


def get_all_worker_infos():
    return sorted(_state["infos"].values(), key=lambda w: w.rank)


def get_current_worker_info():
    return _state["self"]
