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
    # horovodrun-style tuning flags: each one sets the HVD_* variable hvd.init() reads on every rank
    ap.add_argument("-H", "--hosts", default=None, help="'localhost:N' (single host only): N becomes -np when -np is absent")
    ap.add_argument("--fusion-threshold-mb", type=float, default=None, help="gradient bucket size limit (HVD_FUSION_THRESHOLD)")
    ap.add_argument("--cycle-time-ms", type=float, default=None, help="accepted and ignored: there is no negotiation cycle")
    ap.add_argument("--cache-capacity", type=int, default=None, help="accepted and ignored: the bucket plan is static")
    ap.add_argument("--timeline-filename", default=None, help="Chrome-trace timeline written by rank 0 (HVD_TIMELINE)")
    ap.add_argument("--timeline-mark-cycles", action="store_true", help="HVD_TIMELINE_MARK_CYCLES=1")
    ap.add_argument("--autotune", action="store_true", help="measure the allreduce variants at start-up (HVD_AUTOTUNE=1)")
    ap.add_argument("--autotune-log-file", default=None, help="HVD_AUTOTUNE_LOG")
    ap.add_argument("--log-level", default=None, choices=["TRACE", "DEBUG", "INFO", "WARNING", "ERROR", "FATAL"],
                    help="HVD_LOG_LEVEL")
    ap.add_argument("--stall-check-time-seconds", type=float, default=None, help="HVD_STALL_CHECK_TIME_SECONDS")
    ap.add_argument("--allreduce-variant", default=None, choices=["auto", "oneshot", "twoshot", "nvls", "nccl"],
                    help="HVD_ALLREDUCE_VARIANT")
    ap.add_argument("--gloo", action="store_true", help="accepted for horovodrun parity (host control traffic always uses gloo)")
    ap.add_argument("--mpi", action="store_true", help="accepted for horovodrun parity and ignored")
    ap.add_argument("--verbose", action="store_true", help="print the launch plan")
    ap.add_argument("-cb", "--check-build", action="store_true", help="print what this build supports and exit")
    ap.add_argument("command", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    if a.check_build:
        return a
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
    # Like torchrun (and unlike a bare Popen): every rank would otherwise start one OpenMP/MKL worker per core of the
    # host, and N ranks' spinning pools then slow each other's launch loop (measured: 106 vs 86 us/step at 2 ranks).
    omp = str(max(1, min(4, (os.cpu_count() or 1) // max(np_, 1))))
    for r in range(np_):
        env = dict(os.environ)
        if np_ > 1:
            env.setdefault("OMP_NUM_THREADS", omp)
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


def tuning_env(a) -> dict:
    """The HVD_* variables implied by horovodrun-style flags."""
    env = {}
    if a.fusion_threshold_mb is not None:
        env["HVD_FUSION_THRESHOLD"] = str(int(a.fusion_threshold_mb * 1024 * 1024))
    if a.timeline_filename:
        env["HVD_TIMELINE"] = a.timeline_filename
    if a.timeline_mark_cycles:
        env["HVD_TIMELINE_MARK_CYCLES"] = "1"
    if a.autotune:
        env["HVD_AUTOTUNE"] = "1"
    if a.autotune_log_file:
        env["HVD_AUTOTUNE_LOG"] = a.autotune_log_file
    if a.log_level:
        env["HVD_LOG_LEVEL"] = a.log_level
    if a.stall_check_time_seconds is not None:
        env["HVD_STALL_CHECK_TIME_SECONDS"] = str(a.stall_check_time_seconds)
    if a.allreduce_variant:
        env["HVD_ALLREDUCE_VARIANT"] = a.allreduce_variant
    return env


def check_build() -> str:
    import horovod_distributed_example_b200 as hvd
    from .ops import native
    mark = lambda ok: "[X]" if ok else "[ ]"           # noqa: E731
    core, cuda = native.core_lib(required=False) is not None, native.cuda_lib(required=False) is not None
    return "\n".join([
        f"hvd_b200 v{hvd.__version__}:", "", "Available Frameworks:", "    [X] PyTorch (Keras-like front end)", "",
        "Available Controllers:", f"    {mark(hvd.mpi_built())} MPI", f"    {mark(hvd.gloo_built())} Gloo", "",
        "Available Tensor Operations:", f"    {mark(cuda)} peer-memory sm_100a kernels (one-shot / two-shot / NVLS)",
        f"    {mark(hvd.nccl_built())} NCCL (baseline path)", f"    {mark(hvd.gloo_built())} Gloo (CPU)", "",
        "Native libraries:", f"    {mark(core)} libhvd_core.so", f"    {mark(cuda)} libhvd_cuda.so"])


def main(argv=None) -> int:
    a = parse_args(argv)
    if a.check_build:
        print(check_build())
        return 0
    np_ = a.np
    if np_ is None and a.hosts:
        total = 0
        for item in a.hosts.split(","):
            host, _, slots = item.partition(":")
            if host not in ("localhost", "127.0.0.1", socket.gethostname()):
                sys.stderr.write(f"hvdrun: host '{host}' is not this machine; only single-host launch is supported\n")
                return 2
            total += int(slots or 1)
        np_ = total
    if np_ is None:
        try:
            import torch
            np_ = max(torch.cuda.device_count(), 1)
        except Exception:
            np_ = 1
    extra = tuning_env(a)
    for item in a.x:
        if "=" in item:
            k, v = item.split("=", 1)
            extra[k] = v
        elif item in os.environ:
            extra[item] = os.environ[item]
    if a.verbose:
        sys.stderr.write(f"hvdrun: {np_} process(es) on this host, env {extra}, command {a.command}\n")
    return launch(np_, a.command, extra, a.master_port, a.timeout)


if __name__ == "__main__":
    sys.exit(main())
