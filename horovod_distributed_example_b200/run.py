"""``hvdrun`` - single-host launcher: one process per GPU (or per slot on CPU).

Stands in for the reference's launch line (``README.md:57``,
``.ps_project/distributed-keras-sample.yaml:8``)::

    mpirun --allow-run-as-root -np N --hostfile ... -bind-to none -map-by slot \
           -x NCCL_DEBUG=INFO -mca pml ob1 -mca btl ^openib python mnist_keras.py

``python -m horovod_distributed_example_b200.run -np N [-x VAR[=VAL]]... script.py args``
sets the env contract ``hvd.init()`` reads (HVD_RANK/SIZE/LOCAL_RANK/LOCAL_SIZE plus
MASTER_ADDR/PORT), does not bind CPUs (``-bind-to none``), forwards ``-x`` variables, and
tears the whole group down when any rank exits non-zero (mpirun semantics).
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
from typing import List


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def parse_args(argv=None):
    ap = argparse.ArgumentParser(prog="hvdrun", description=__doc__.split("\n")[0])
    ap.add_argument("-np", "--num-proc", type=int, default=None, dest="np",
                    help="number of processes (default: number of visible GPUs, else 1)")
    ap.add_argument("-x", action="append", default=[], metavar="VAR[=VAL]",
                    help="export an environment variable to every rank (as mpirun -x)")
    ap.add_argument("--master-port", type=int, default=0)
    ap.add_argument("--allow-run-as-root", action="store_true", help="accepted for mpirun parity")
    ap.add_argument("--hostfile", default=None, help="accepted; only single-host launch is supported")
    ap.add_argument("-bind-to", dest="bind_to", default="none")
    ap.add_argument("-map-by", dest="map_by", default="slot")
    ap.add_argument("-mca", nargs=2, action="append", default=[], help="accepted and ignored")
    ap.add_argument("--timeout", type=float, default=0.0, help="kill the job after this many seconds")
    ap.add_argument("command", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    if not a.command:
        ap.error("no command given")
    if a.command[0] == "--":
        a.command = a.command[1:]
    return a


def launch(np_: int, command: List[str], extra_env=None, master_port: int = 0, timeout: float = 0.0) -> int:
    port = master_port or _free_port()
    procs: List[subprocess.Popen] = []
    cmd = command
    if cmd[0].endswith(".py"):
        cmd = [sys.executable] + cmd
    for r in range(np_):
        env = dict(os.environ)
        env.update(extra_env or {})
        env.update({"HVD_RANK": str(r), "HVD_SIZE": str(np_), "HVD_LOCAL_RANK": str(r),
                    "HVD_LOCAL_SIZE": str(np_), "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
        procs.append(subprocess.Popen(cmd, env=env, start_new_session=True))
    t0 = time.time()
    rc = 0
    try:
        alive = set(range(np_))
        while alive:
            for i in list(alive):
                code = procs[i].poll()
                if code is None:
                    continue
                alive.discard(i)
                if code != 0 and rc == 0:
                    rc = code
                    sys.stderr.write(f"hvdrun: rank {i} exited with code {code}; terminating the job\n")
                    _kill_all(procs)
            if timeout and time.time() - t0 > timeout and alive:
                sys.stderr.write("hvdrun: timeout; terminating the job\n")
                rc = rc or 124
                _kill_all(procs)
            time.sleep(0.05)
    except KeyboardInterrupt:
        _kill_all(procs)
        rc = 130
    return rc


def _kill_all(procs):
    for p in procs:
        if p.poll() is None:
            try:
                os.killpg(p.pid, signal.SIGTERM)   # exact process group we started
            except ProcessLookupError:
                pass
    deadline = time.time() + 5
    for p in procs:
        try:
            p.wait(max(0.0, deadline - time.time()))
        except subprocess.TimeoutExpired:
            try:
                os.killpg(p.pid, signal.SIGKILL)
            except ProcessLookupError:
                pass


def main(argv=None) -> int:
    a = parse_args(argv)
    np_ = a.np
    if np_ is None:
        try:
            import torch
            np_ = max(torch.cuda.device_count(), 1)
        except Exception:
            np_ = 1
    extra = {}
    for item in a.x:
        if "=" in item:
            k, v = item.split("=", 1)
            extra[k] = v
        elif item in os.environ:
            extra[item] = os.environ[item]
    return launch(np_, a.command, extra, a.master_port, a.timeout)


if __name__ == "__main__":
    sys.exit(main())
