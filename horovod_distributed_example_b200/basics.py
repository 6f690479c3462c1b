"""Process-level state: ``init / shutdown / size / rank / local_rank`` and friends.

Capability parity with the reference's use of ``hvd.init()`` (``mnist_keras.py:30``,
``tensorflow2_keras_mnist.py:25``), ``hvd.size()`` (``mnist_keras.py:42,84``),
``hvd.rank()`` (``mnist_keras.py:101,111,116``) and ``hvd.local_rank()``
(``mnist_keras.py:35``, ``tensorflow2_keras_mnist.py:32``).

Design differences from Horovod 0.18.1:

* no background coordination thread and no MPI - one process per GPU, bootstrap over
  ``torch.distributed`` (NCCL for device tensors, gloo for host tensors/control);
* ``init()`` pins the GPU itself (``torch.cuda.set_device(local_rank)``) so the
  reference's explicit ``visible_device_list`` dance is a no-op here;
* with no launcher at all ``init()`` yields a size-1 world (the reference's YAML
  actually runs the script outside mpirun - SURVEY.md section 2.1 quirks).
"""
from __future__ import annotations

import atexit
import datetime
import logging
import os
import socket
from dataclasses import dataclass, field
from typing import Optional

from .config import Config

log = logging.getLogger("hvd")


@dataclass
class _State:
    initialized: bool = False
    rank: int = 0
    size: int = 1
    local_rank: int = 0
    local_size: int = 1
    cross_rank: int = 0
    cross_size: int = 1
    device: Optional[object] = None      # torch.device
    config: Optional[Config] = None
    owns_process_group: bool = False
    cpu_group: Optional[object] = None   # gloo group for host tensors
    peer: Optional[object] = None        # parallel.peer.PeerContext (lazily created)
    timeline: Optional[object] = None
    step: int = 0
    shutdown_hooks: list = field(default_factory=list)     # run first by shutdown() (e.g. the engine's timeline drain)


_state = _State()

# (rank, size, local_rank, local_size) env spellings, in priority order.  The last
# two are what the reference's launch line provides (Open MPI, README.md:57) and
# what torchrun / the driver's bench launcher provides.
_ENV_CONTRACTS = (
    ("HVD_RANK", "HVD_SIZE", "HVD_LOCAL_RANK", "HVD_LOCAL_SIZE"),
    ("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK",
     "OMPI_COMM_WORLD_LOCAL_SIZE"),
    ("PMI_RANK", "PMI_SIZE", "MPI_LOCALRANKID", "MPI_LOCALNRANKS"),
    ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE"),
)


def discover_env(env=None):
    """Return (rank, size, local_rank, local_size) from the launcher's environment.

    Falls back to a singleton world when no launcher variables are present.
    """
    if env is None:
        # the process environment: the native runtime core owns the contract table (csrc/core/runtime.cpp,
        # hvd_core_discover_env); the loop below is its twin for explicit dicts and hosts without the library
        core = _native_core()
        if core is not None:
            import ctypes
            v = (ctypes.c_int * 4)()
            if core.hvd_core_discover_env(v, None) != 0:
                raise ValueError("launcher environment is inconsistent: rank outside [0, size)")
            return int(v[0]), int(v[1]), int(v[2]), int(v[3])
        env = os.environ
    for rk, sz, lr, ls in _ENV_CONTRACTS:
        if rk in env and sz in env:
            rank, size = int(env[rk]), int(env[sz])
            local_rank = int(env.get(lr, rank))
            local_size = int(env.get(ls, size))
            if not (0 <= rank < size):
                raise ValueError(f"{rk}={rank} out of range for {sz}={size}")
            return rank, size, local_rank, local_size
    return 0, 1, 0, 1


def _native_core():
    try:
        from .ops import native
        return native.core_lib(required=False)
    except Exception:  # pragma: no cover - no toolchain and no prebuilt library
        return None


def _setup_logging(level: str):
    lvl = getattr(logging, level, logging.WARNING)
    if not log.handlers:
        h = logging.StreamHandler()
        h.setFormatter(logging.Formatter("[hvd %(levelname)s r%(rank)s] %(message)s"))
        log.addHandler(h)
        log.propagate = False

        class _RankFilter(logging.Filter):
            def filter(self, record):
                record.rank = _state.rank
                return True
        log.addFilter(_RankFilter())
    log.setLevel(lvl)


def init(comm=None, *, backend: Optional[str] = None, device: Optional[str] = None) -> None:
    """Initialise the framework.  Idempotent; works with no launcher (size 1).

    ``comm`` is accepted for Horovod signature compatibility and ignored.
    ``backend``/``device`` override auto-detection (tests use ``device='cpu'``).
    """
    import torch

    if _state.initialized:
        return
    cfg = Config.from_env()
    core = _native_core()
    if core is not None:
        # init state lives in the native runtime core (Horovod: operations.cc horovod_init / rank / size); C and
        # C++ callers of the C ABI read the same values through hvd_core_rank() etc.
        if core.hvd_core_init() != 0:
            raise ValueError("launcher environment is inconsistent: rank outside [0, size)")
        rank, size = core.hvd_core_rank(), core.hvd_core_size()
        local_rank, local_size = core.hvd_core_local_rank(), core.hvd_core_local_size()
    else:
        rank, size, local_rank, local_size = discover_env()
    _state.rank, _state.size = rank, size
    _state.local_rank, _state.local_size = local_rank, local_size
    _state.cross_rank = rank // max(local_size, 1)
    _state.cross_size = max(size // max(local_size, 1), 1)
    _state.config = cfg
    _setup_logging(cfg.log_level)
    from .utils import nvtx
    nvtx.configure(cfg.nvtx, rank=rank, level=cfg.log_level)

    want_cuda = torch.cuda.is_available() if device is None else device.startswith("cuda")
    if want_cuda:
        ndev = torch.cuda.device_count()
        if local_rank >= ndev:
            raise RuntimeError(
                f"local_rank {local_rank} but only {ndev} CUDA device(s) visible: "
                "one process per GPU is required")
        torch.cuda.set_device(local_rank)
        _state.device = torch.device("cuda", local_rank)
    else:
        _state.device = torch.device("cpu")

    if size > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            if backend is None:
                backend = "cpu:gloo,cuda:nccl" if want_cuda else "gloo"
            kw = {}
            if want_cuda and "nccl" in backend:
                kw["device_id"] = _state.device
            dist.init_process_group(
                backend=backend, rank=rank, world_size=size,
                timeout=datetime.timedelta(seconds=max(60.0, cfg.stall_check_time * 5)), **kw)
            _state.owns_process_group = True
        else:
            if dist.get_rank() != rank or dist.get_world_size() != size:
                raise RuntimeError("torch.distributed is already initialised with a different world")
        # host-side control traffic (barriers, object exchange, fd bootstrap) always goes over a
        # dedicated gloo group so it can never interleave with device collectives
        if str(dist.get_backend()) == "gloo":
            _state.cpu_group = dist.group.WORLD
        else:
            _state.cpu_group = dist.new_group(backend="gloo")

    if cfg.timeline and rank == 0:
        from .utils.timeline import Timeline
        _state.timeline = Timeline(cfg.timeline, mark_cycles=cfg.timeline_mark_cycles)

    _state.initialized = True
    atexit.register(shutdown)
    log.info("init: rank %d/%d local_rank %d host %s device %s",
             rank, size, local_rank, socket.gethostname(), _state.device)


def shutdown() -> None:
    """Tear down peer memory, the timeline and (if we created it) the process group."""
    if not _state.initialized:
        return
    for hook in list(_state.shutdown_hooks):
        try:
            hook()
        except Exception as e:  # pragma: no cover - best effort at exit
            log.warning("shutdown hook failed: %s: %s", type(e).__name__, e)
    _state.shutdown_hooks.clear()
    try:
        if _state.peer is not None:
            _state.peer.close()
    except Exception as e:  # pragma: no cover - best effort at exit
        log.debug("peer close failed: %s", e)
    _state.peer = None
    import sys as _sys
    _peer_mod = _sys.modules.get(__package__ + ".parallel.peer")
    if _peer_mod is not None:
        _peer_mod.reset()                # a later init() builds a fresh context (fresh epochs, pads, buffers)
    if _state.timeline is not None:
        _state.timeline.close()
        _state.timeline = None
    if _state.size > 1 and _state.owns_process_group:
        import torch.distributed as dist
        if dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception as e:  # pragma: no cover
                log.debug("destroy_process_group failed: %s", e)
    _state.__init__()
    core = _native_core()
    if core is not None:
        core.hvd_core_shutdown()


def _require_init():
    if not _state.initialized:
        # Horovod raises ValueError from size()/rank() before init.
        raise ValueError("Horovod has not been initialized; use hvd.init().")


def is_initialized() -> bool:
    return _state.initialized


def size() -> int:
    _require_init()
    return _state.size


def rank() -> int:
    _require_init()
    return _state.rank


def local_rank() -> int:
    _require_init()
    return _state.local_rank


def local_size() -> int:
    _require_init()
    return _state.local_size


def cross_rank() -> int:
    _require_init()
    return _state.cross_rank


def cross_size() -> int:
    _require_init()
    return _state.cross_size


def device():
    """The torch device this process computes on (``cuda:<local_rank>`` or ``cpu``)."""
    _require_init()
    return _state.device


def config() -> Config:
    _require_init()
    return _state.config


def state() -> _State:
    return _state


# Build-capability probes kept for Horovod API compatibility.
def mpi_enabled() -> bool:
    return False


def mpi_built() -> bool:
    return False


def gloo_enabled() -> bool:
    return True


def gloo_built() -> bool:
    return True


def nccl_built() -> bool:
    import torch.distributed as dist
    return dist.is_nccl_available()


def mpi_threads_supported() -> bool:
    return False
