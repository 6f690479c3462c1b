import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with `pytest -m gpu` under gpurun)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 GPUs")


@pytest.fixture
def hvd_single(monkeypatch):
    """A fresh size-1 hvd world with launcher env removed."""
    for k in list(os.environ):
        if k.startswith(("HVD_", "HOROVOD_", "OMPI_COMM_WORLD_", "PMI_")) or k in (
                "RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
            monkeypatch.delenv(k, raising=False)
    import horovod_distributed_example_b200 as hvd
    hvd.shutdown()
    hvd.init(device="cpu")
    yield hvd
    hvd.shutdown()


def run_workers(fn_name, world_size=2, timeout=240, env=None, module=None):
    """Spawn `world_size` CPU/gloo ranks running tests/_workers.py::<fn_name>; return rcs."""
    import socket
    import subprocess
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    procs = []
    for r in range(world_size):
        e = dict(os.environ)
        e.update({"HVD_RANK": str(r), "HVD_SIZE": str(world_size), "HVD_LOCAL_RANK": str(r),
                  "HVD_LOCAL_SIZE": str(world_size), "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
                  "CUDA_VISIBLE_DEVICES": "", "OMP_NUM_THREADS": "2"})
        e.update(env or {})
        procs.append(subprocess.Popen(
            [sys.executable, os.path.join(ROOT, "tests", module or "_workers.py"), fn_name],
            env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append((p.returncode, out))
    return outs
