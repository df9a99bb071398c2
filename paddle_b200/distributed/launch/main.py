"""Collective launcher (reference: python/paddle/distributed/launch/main.py:20-296, context/args_envs.py:57-225,
controllers/{controller,collective,master}.py, job/{pod,container}.py, utils/kv_server.py).

One process per device is started with the PADDLE_* environment contract (Appendix A "Worker env contract"),
stdout/stderr go to ``{log_dir}/workerlog.N``.  Multi-node: node 0 runs an HTTP KV master on ``--master``; every
node registers its candidate endpoints under ``/job/info`` and waits until ``--nnodes`` peers are present, then
ranks are assigned by sorted order.  The watch loop polls every 2 s; on a failed worker it tails the log, stops the
pod and - with ``--max_restart`` > 0 and ``--elastic_level`` >= 0 - relaunches the whole pod (workers are expected
to resume from their last checkpoint).
"""
from __future__ import annotations

import argparse
import json
import shutil
import os
import signal
import socket
import subprocess
import sys
import threading
import time
import urllib.request
from http.server import BaseHTTPRequestHandler, HTTPServer

_ENV_ARGS = {
    "PADDLE_MASTER": "master", "PADDLE_DEVICES": "devices", "PADDLE_NNODES": "nnodes", "PADDLE_RUN_MODE": "run_mode",
    "PADDLE_LOG_LEVEL": "log_level", "PADDLE_LOG_OVERWRITE": "log_overwrite", "PADDLE_SORT_IP": "sort_ip",
    "PADDLE_NPROC_PER_NODE": "nproc_per_node", "PADDLE_JOB_ID": "job_id", "PADDLE_RANK": "rank",
    "PADDLE_LOG_DIR": "log_dir", "PADDLE_MAX_RESTART": "max_restart", "PADDLE_ELASTIC_LEVEL": "elastic_level",
    "PADDLE_ELASTIC_TIMEOUT": "elastic_timeout", "PADDLE_SERVER_NUM": "server_num", "PADDLE_TRAINER_NUM": "trainer_num",
    "PADDLE_WITH_GLOO": "with_gloo", "PADDLE_START_PORT": "start_port", "PADDLE_IPS": "ips",
}


def parse_args(argv=None):
    p = argparse.ArgumentParser(prog="python -m paddle.distributed.launch")
    b = p.add_argument_group("Base Parameters")
    b.add_argument("--master", type=str, default=None, help="the master/rendezvous server, ip:port")
    b.add_argument("--legacy", type=lambda s: s.lower() in ("1", "true"), default=False)
    b.add_argument("--rank", type=int, default=-1, help="the node rank")
    b.add_argument("--log_level", type=str, default="INFO")
    b.add_argument("--log_overwrite", type=lambda s: s.lower() in ("1", "true"), default=False)
    b.add_argument("--sort_ip", type=lambda s: s.lower() in ("1", "true"), default=False)
    b.add_argument("--enable_gpu_log", type=lambda s: s.lower() in ("1", "true"), default=True)
    b.add_argument("--nnodes", type=str, default="1", help="the number of nodes, e.g. 4 or 2:4 (elastic range)")
    b.add_argument("--nproc_per_node", type=int, default=None)
    b.add_argument("--log_dir", type=str, default="log")
    b.add_argument("--run_mode", type=str, default=None, help="collective (default), ps, rpc")
    b.add_argument("--job_id", type=str, default="default")
    b.add_argument("--devices", "--gpus", "--xpus", type=str, default=None, dest="devices")
    b.add_argument("--host", type=str, default=None)
    b.add_argument("--ips", type=str, default=None, help="[DEPRECATED] node ips")
    b.add_argument("--start_port", type=int, default=6070)
    b.add_argument("--auto_parallel_config", type=str, default=None)
    b.add_argument("--auto_tuner_json", type=str, default=None)
    b.add_argument("training_script", type=str)
    b.add_argument("training_script_args", nargs=argparse.REMAINDER)
    ps = p.add_argument_group("Parameter-Server Parameters")
    ps.add_argument("--servers", type=str, default="")
    ps.add_argument("--trainers", type=str, default="")
    ps.add_argument("--trainer_num", type=int, default=None)
    ps.add_argument("--server_num", type=int, default=None)
    ps.add_argument("--gloo_port", type=int, default=6767)
    ps.add_argument("--with_gloo", type=str, default="1")
    e = p.add_argument_group("Elastic Parameters")
    e.add_argument("--max_restart", type=int, default=3)
    e.add_argument("--elastic_level", type=int, default=-1)
    e.add_argument("--elastic_timeout", type=int, default=30)
    args = p.parse_args(argv)
    for env, name in _ENV_ARGS.items():
        if env in os.environ and getattr(args, name, None) in (None, p.get_default(name)):
            cur = getattr(args, name, None)
            val = os.environ[env]
            setattr(args, name, type(cur)(val) if cur is not None and not isinstance(cur, bool) else val)
    return args


def _free_ports(n, start=None):
    ports, socks = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("", 0))
        ports.append(s.getsockname()[1])
        socks.append(s)
    for s in socks:
        s.close()
    return ports


def _host_ip(args):
    if args.host:
        return args.host
    try:
        return socket.gethostbyname(socket.gethostname())
    except OSError:
        return "127.0.0.1"


def _detect_devices(args):
    if args.devices:
        return [d for d in str(args.devices).split(",") if d != ""]
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        return [str(i) for i in range(len(vis.split(",")))]
    try:
        import torch
        n = torch.cuda.device_count()
    except Exception:  # noqa: BLE001
        n = 0
    return [str(i) for i in range(n)]


# ---------------------------------------------------------------------------------------------- HTTP KV master
class _KVHandler(BaseHTTPRequestHandler):
    store: dict = {}
    lock = threading.Lock()

    def log_message(self, *a):
        pass

    def do_GET(self):
        with self.lock:
            prefix = self.path
            data = {k: v for k, v in self.store.items() if k.startswith(prefix)}
        body = json.dumps(data).encode()
        self.send_response(200)
        self.send_header("Content-Length", str(len(body)))
        self.end_headers()
        self.wfile.write(body)

    def do_PUT(self):
        n = int(self.headers.get("Content-Length", 0))
        val = self.rfile.read(n).decode()
        with self.lock:
            self.store[self.path] = val
        self.send_response(200)
        self.end_headers()

    do_POST = do_PUT

    def do_DELETE(self):
        with self.lock:
            for k in [k for k in self.store if k.startswith(self.path)]:
                del self.store[k]
        self.send_response(200)
        self.end_headers()


class KVServer:
    def __init__(self, port):
        self.httpd = HTTPServer(("", port), _KVHandler)
        self.thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)

    def start(self):
        self.thread.start()

    def stop(self):
        self.httpd.shutdown()
        self.httpd.server_close()


class KVClient:
    def __init__(self, endpoint):
        self.base = endpoint if endpoint.startswith("http") else "http://" + endpoint

    def put(self, key, value):
        req = urllib.request.Request(self.base + key, data=value.encode(), method="PUT")
        try:
            urllib.request.urlopen(req, timeout=3).read()
            return True
        except Exception:  # noqa: BLE001
            return False

    def get_prefix(self, key):
        try:
            return json.loads(urllib.request.urlopen(self.base + key, timeout=3).read().decode())
        except Exception:  # noqa: BLE001
            return {}

    def delete(self, key):
        try:
            urllib.request.urlopen(urllib.request.Request(self.base + key, method="DELETE"), timeout=3).read()
        except Exception:  # noqa: BLE001
            pass


class HTTPMaster:
    """sync_peers over the KV server (reference: controllers/master.py:73-183)."""

    def __init__(self, args, host_ip):
        self.endpoint = args.master
        self.server = None
        self.role_main = False
        host, port = self.endpoint.rsplit(":", 1)
        if args.rank in (0, -1) and host in (host_ip, "127.0.0.1", "localhost", socket.gethostname()):
            try:
                self.server = KVServer(int(port))
                self.server.start()
                self.role_main = True
            except OSError:
                self.server = None
        self.client = KVClient(self.endpoint)

    def sync_peers(self, prefix, key, value, size, rank=-1, timeout=600):
        if size < 2:
            return [value], 0
        k = f"{prefix}/{key}" if rank < 0 else f"{prefix}/{rank:08d}-{key}"
        t0 = time.time()
        while not self.client.put(k, value):
            if time.time() - t0 > timeout:
                raise TimeoutError(f"cannot reach master {self.endpoint}")
            time.sleep(0.5)
        while True:
            peers = self.client.get_prefix(prefix + "/")
            if len(peers) == size:
                keys = sorted(peers)
                return [peers[x] for x in keys], keys.index(k)
            if time.time() - t0 > timeout:
                raise TimeoutError(f"only {len(peers)}/{size} peers joined before timeout")
            time.sleep(0.5)

    def stop(self):
        if self.server is not None:
            self.server.stop()


# ---------------------------------------------------------------------------------------------- pod / containers
class Container:
    def __init__(self, entrypoint, env, log_path, overwrite):
        self.entrypoint, self.env, self.log_path, self.overwrite = entrypoint, env, log_path, overwrite
        self.proc = None
        self._log = None

    def start(self):
        os.makedirs(os.path.dirname(self.log_path) or ".", exist_ok=True)
        self._log = open(self.log_path, "w" if self.overwrite else "a")
        env = dict(os.environ)
        env.update(self.env)
        self.proc = subprocess.Popen(self.entrypoint, env=env, stdout=self._log, stderr=subprocess.STDOUT,
                                     start_new_session=True)

    def status(self):
        if self.proc is None:
            return "init"
        rc = self.proc.poll()
        if rc is None:
            return "running"
        return "completed" if rc == 0 else "failed"

    def exit_code(self):
        return self.proc.poll() if self.proc else None

    def terminate(self, force=False):
        if self.proc is not None and self.proc.poll() is None:
            try:
                os.killpg(self.proc.pid, signal.SIGKILL if force else signal.SIGTERM)
            except ProcessLookupError:
                pass
        if self._log:
            self._log.flush()

    def wait(self, timeout=None):
        try:
            self.proc.wait(timeout)
        except subprocess.TimeoutExpired:
            pass

    def tail(self, n=2000):
        try:
            with open(self.log_path, "rb") as f:
                f.seek(0, 2)
                size = f.tell()
                f.seek(max(size - n, 0))
                return f.read().decode(errors="replace")
        except OSError:
            return ""


class Pod:
    def __init__(self):
        self.containers = []

    def deploy(self):
        for c in self.containers:
            c.start()

    def status(self):
        st = [c.status() for c in self.containers]
        if any(s == "failed" for s in st):
            return "failed"
        if all(s == "completed" for s in st):
            return "completed"
        return "running"

    def failed(self):
        return [c for c in self.containers if c.status() == "failed"]

    def stop(self, grace=10):
        for c in self.containers:
            c.terminate()
        t0 = time.time()
        for c in self.containers:
            c.wait(max(0.1, grace - (time.time() - t0)))
        for c in self.containers:
            c.terminate(force=True)


class Watcher:
    """GPU utilisation logger (reference: python/paddle/distributed/launch/job/watcher? -> launch/controllers/watcher.py):
    one ``nvidia-smi`` query every ``interval`` seconds appended to ``<log_dir>/<job_id>.gpu.log`` while the pod runs."""
    QUERY = "index,utilization.gpu,memory.used,memory.total,power.draw,clocks.sm"

    def __init__(self, args, devices, interval=10):
        self._stop = threading.Event()
        self._thread = None
        self.path = os.path.join(args.log_dir, f"{args.job_id}.gpu.log")
        self._devices = devices
        self._interval = interval
        self._smi = shutil.which("nvidia-smi")

    def start(self):
        if not self._smi or not self._devices:
            return self
        os.makedirs(os.path.dirname(self.path) or ".", exist_ok=True)
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def _run(self):
        with open(self.path, "a") as f:
            f.write("time," + self.QUERY + "\n")
            while not self._stop.is_set():
                try:
                    out = subprocess.run([self._smi, f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-i", ",".join(self._devices)], capture_output=True, text=True, timeout=10).stdout
                    now = time.strftime("%Y-%m-%d %H:%M:%S")
                    for line in out.strip().splitlines():
                        f.write(f"{now},{line.strip()}\n")
                    f.flush()
                except Exception:  # noqa: BLE001  (the logger must never take the job down)
                    pass
                self._stop.wait(self._interval)

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=5)


def run_auto_tuner(args):
    """``--auto_tuner_json cfg.json``: search the parallel-strategy space by launching the training script once per
    candidate (reference: launch/main.py auto-tuner branch + distributed/auto_tuner).  The script reads its candidate
    from ``PADDLE_AUTO_TUNER_CFG`` and prints one JSON line with the metric named in ``metric_cfg.name``."""
    from paddle_b200.distributed.auto_tuner import AutoTuner
    from paddle_b200.distributed.auto_tuner.tuner import launch_trial
    with open(args.auto_tuner_json) as f:
        cfg = json.load(f)
    devices = _detect_devices(args)
    nproc = args.nproc_per_node or (len(devices) if devices else 1)
    cfg.setdefault("num_gpus", nproc)
    tuner = AutoTuner(cfg)
    best = tuner.tune(lambda c: launch_trial(args.training_script, c, nproc, os.path.join(args.log_dir, "auto_tuner"),
                                             args.training_script_args, timeout=cfg.get("max_time_per_task", 3600)))
    os.makedirs(args.log_dir, exist_ok=True)
    tuner.recorder.store_history(os.path.join(args.log_dir, "auto_tuner_history.csv"))
    sys.stdout.write("AUTO_TUNER best: " + json.dumps(best) + "\n")
    return 0 if best is not None else 1


def build_pod(args, node_rank, nnodes, all_endpoints, local_range, devices, host_ip, master_ep, restart=0):
    pod = Pod()
    global_size = len(all_endpoints)
    lo, hi = local_range
    for i, gr in enumerate(range(lo, hi)):
        env = {
            "PADDLE_MASTER": master_ep, "PADDLE_GLOBAL_SIZE": str(global_size), "PADDLE_LOCAL_SIZE": str(hi - lo),
            "PADDLE_GLOBAL_RANK": str(gr), "PADDLE_LOCAL_RANK": str(i), "PADDLE_NNODES": str(nnodes),
            "PADDLE_CURRENT_ENDPOINT": all_endpoints[gr], "PADDLE_TRAINER_ID": str(gr),
            "PADDLE_TRAINERS_NUM": str(global_size), "PADDLE_RANK_IN_NODE": str(i),
            "PADDLE_TRAINER_ENDPOINTS": ",".join(all_endpoints), "PADDLE_LOG_DIR": args.log_dir,
            "POD_NAME": f"{args.job_id}-{node_rank}", "PADDLE_JOB_ID": args.job_id, "PADDLE_RESTART_COUNT": str(restart),
        }
        if devices:
            env["FLAGS_selected_gpus"] = devices[i % len(devices)]
        else:
            env["PADDLE_DISTRI_BACKEND"] = "gloo"
        entry = [sys.executable, "-u", args.training_script, *args.training_script_args]
        pod.containers.append(Container(entry, env, os.path.join(args.log_dir, f"workerlog.{gr if nnodes == 1 else i}"),
                                        args.log_overwrite or restart == 0))
    return pod


def launch(argv=None):
    args = parse_args(argv)
    if args.run_mode not in (None, "collective"):
        raise NotImplementedError(f"run_mode={args.run_mode}: only collective mode is in scope (PS/RPC/IPU are not)")
    if args.auto_tuner_json:
        return run_auto_tuner(args)
    devices = _detect_devices(args)
    nproc = args.nproc_per_node or (len(devices) if devices else 1)
    if devices and nproc > len(devices) and args.devices:
        raise ValueError(f"nproc_per_node {nproc} exceeds the {len(devices)} visible devices")
    nnodes = int(str(args.nnodes).split(":")[0])
    host_ip = _host_ip(args) if nnodes > 1 else "127.0.0.1"
    master = None
    restart = 0
    stop_flag = {"stop": False}
    pod_holder = {}

    def on_signal(signum, frame):
        stop_flag["stop"] = True
        if "pod" in pod_holder:
            pod_holder["pod"].stop(grace=5)
        sys.exit(128 + signum)
    for sg in (signal.SIGTERM, signal.SIGINT, signal.SIGABRT):
        signal.signal(sg, on_signal)

    while True:
        ports = _free_ports(nproc)
        my_eps = [f"{host_ip}:{p}" for p in ports]
        if nnodes > 1:
            if not args.master:
                raise ValueError("--master ip:port is required when --nnodes > 1")
            master = master or HTTPMaster(args, host_ip)
            info = json.dumps({"name": f"{args.job_id}-{host_ip}-{os.getpid()}", "endpoints": my_eps, "restart": restart})
            peers, node_rank = master.sync_peers(f"/{args.job_id}/info/{restart}", f"{host_ip}-{os.getpid()}", info, nnodes,
                                                 rank=args.rank)
            all_eps, ranges, off = [], [], 0
            for pj in peers:
                eps = json.loads(pj)["endpoints"]
                ranges.append((off, off + len(eps)))
                all_eps += eps
                off += len(eps)
            local_range = ranges[node_rank]
            master_ep = all_eps[0]
        else:
            node_rank, all_eps, local_range = 0, my_eps, (0, nproc)
            master_ep = args.master or my_eps[0]
        pod = build_pod(args, node_rank, nnodes, all_eps, local_range, devices, host_ip, master_ep, restart)
        pod_holder["pod"] = pod
        pod.deploy()
        watcher = Watcher(args, devices).start() if args.enable_gpu_log else None
        status = "running"
        while status == "running" and not stop_flag["stop"]:
            time.sleep(2 if nproc > 0 else 0.5)
            status = pod.status()
        if watcher is not None:
            watcher.stop()
        if status == "completed":
            if master:
                master.stop()
            return 0
        failed = pod.failed()
        for c in failed[:1]:
            sys.stderr.write(f"LAUNCH: worker failed with exit code {c.exit_code()}, log tail of {c.log_path}:\n")
            sys.stderr.write(c.tail() + "\n")
        pod.stop()
        if args.elastic_level >= 0 and restart < args.max_restart:
            restart += 1
            sys.stderr.write(f"LAUNCH: restarting pod ({restart}/{args.max_restart})\n")
            continue
        if master:
            master.stop()
        sys.exit(failed[0].exit_code() if failed and failed[0].exit_code() else 1)


if __name__ == "__main__":
    launch()
