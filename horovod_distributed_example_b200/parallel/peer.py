"""Peer-mapped symmetric memory: bootstrap, allocation and the tensor-level collectives.

Stands in for Horovod's communicator setup (``mpi_context.cc`` / NCCL unique-id broadcast,
SURVEY.md K-C4) and its fusion buffer.  One ``PeerContext`` per process:

* allocates symmetric buffers - the same size on every rank, each mapped into every
  peer's address space - with the CUDA VMM driver API (``csrc/core/symm.cpp``; POSIX fds
  travel over a unix socket with SCM_RIGHTS) or, as a fallback, ``cudaMalloc`` +
  CUDA IPC handles;
* binds an NVLS multicast object over each buffer where the box exposes it;
* owns the signal pads / epoch counters used by the kernels in ``csrc/comm``;
* offers ``allreduce_`` / ``broadcast_`` for arbitrary CUDA tensors by staging through a
  small symmetric scratch buffer (metric vectors, non-arena tensors).  Arena gradients do
  not go through the staging path - ``parallel/fused.py`` reduces them in place.

An *emulated* context (``PeerContext.emulated(n)``) puts n virtual ranks on ONE GPU and runs
each collective as a single cooperative launch whose ``blockIdx.y`` is the rank.  It exists so
the multi-rank kernels can be tested on a single-GPU box; separate waiting launches on one
GPU are not safe (B200_PROFILING.md).
"""
from __future__ import annotations

import ctypes
import os
import tempfile
from typing import List, Optional

import torch

from .. import basics
from ..ops import native

MAX_RANKS = 8
MAX_CTAS = 320
NUM_SLOTS = 32

# slots of the signal pad
SLOT_STAGING = 30
SLOT_BCAST = 31

_ctx: Optional["PeerContext"] = None
_unavailable_reason: Optional[str] = None


class HvdCommCtx(ctypes.Structure):
    _fields_ = [("rank", ctypes.c_int), ("nranks", ctypes.c_int),
                ("base", ctypes.c_void_p * MAX_RANKS), ("mc_base", ctypes.c_void_p),
                ("pad", ctypes.c_void_p * MAX_RANKS), ("epoch", ctypes.c_void_p * MAX_RANKS),
                ("error", ctypes.c_void_p), ("timeout_ns", ctypes.c_ulonglong)]


class HvdFusedArgs(ctypes.Structure):
    _fields_ = [("start", ctypes.c_longlong), ("end", ctypes.c_longlong),
                ("grad_off", ctypes.c_longlong), ("param_off", ctypes.c_longlong),
                ("slot0_off", ctypes.c_longlong), ("slot1_off", ctypes.c_longlong),
                ("out_off", ctypes.c_longlong), ("hyper", ctypes.c_void_p),
                ("scale", ctypes.c_float), ("kind", ctypes.c_int), ("variant", ctypes.c_int),
                ("slot", ctypes.c_int), ("exit_barrier", ctypes.c_int)]


class HvdLLArgs(ctypes.Structure):
    _fields_ = [("in_off", ctypes.c_longlong), ("out_off", ctypes.c_longlong), ("n", ctypes.c_longlong),
                ("scale", ctypes.c_float), ("ll", ctypes.c_void_p * MAX_RANKS), ("cap", ctypes.c_longlong)]


ONESHOT, TWOSHOT, NVLS, LL = 0, 1, 2, 3
VARIANT_NAMES = {ONESHOT: "oneshot", TWOSHOT: "twoshot", NVLS: "nvls", LL: "ll"}
#: messages up to this size take the barrier-free LL one-shot kernel (packets of 2 floats + 2 flags)
LL_MAX_BYTES = int(os.environ.get("HVD_LL_MAX_BYTES", 128 << 10))


class _DevPtr:
    """Expose a raw device pointer to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4",
                                         "data": (int(ptr), False), "version": 3, "strides": None}


def tensor_from_ptr(ptr: int, nbytes: int, device, dtype=torch.float32) -> torch.Tensor:
    t = torch.as_tensor(_DevPtr(ptr, nbytes), device=device)
    return t if dtype == torch.float32 else t.view(dtype)


class SymmBuffer:
    """One symmetric allocation: local pointer, every peer's mapping of it, optional multicast."""

    def __init__(self, nbytes: int, ptrs: List[int], mc_ptr: int, local_rank_index: int, device):
        self.nbytes = nbytes
        self.ptrs = ptrs                    # ptrs[r] = address of rank r's copy in THIS process
        self.mc_ptr = mc_ptr                # 0 when NVLS is unavailable
        self.local_index = local_rank_index
        self.device = device
        self.tensor = tensor_from_ptr(ptrs[local_rank_index], nbytes, device)
        self._keep = []

    def peer_tensor(self, r: int) -> torch.Tensor:
        return tensor_from_ptr(self.ptrs[r], self.nbytes, self.device)


def _declare(lib):
    if getattr(lib, "_hvd_comm_declared", False):
        return
    P = ctypes.POINTER
    lib.hvd_fused_allreduce_launch.restype = ctypes.c_int
    lib.hvd_fused_allreduce_launch.argtypes = [P(HvdCommCtx), P(HvdFusedArgs), ctypes.c_int, ctypes.c_int,
                                               ctypes.c_void_p, ctypes.c_int]
    lib.hvd_ll_allreduce_launch.restype = ctypes.c_int
    lib.hvd_ll_allreduce_launch.argtypes = [P(HvdCommCtx), P(HvdLLArgs), ctypes.c_void_p, ctypes.c_int]
    lib.hvd_bcast_launch.restype = ctypes.c_int
    lib.hvd_bcast_launch.argtypes = [P(HvdCommCtx), ctypes.c_int, ctypes.c_int, ctypes.c_longlong,
                                     ctypes.c_longlong, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    lib.hvd_cuda_malloc.restype = ctypes.c_int
    lib.hvd_cuda_malloc.argtypes = [P(ctypes.c_void_p), ctypes.c_longlong]
    lib.hvd_cuda_free.argtypes = [ctypes.c_void_p]
    lib.hvd_ipc_get_handle.restype = ctypes.c_int
    lib.hvd_ipc_get_handle.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.hvd_ipc_open.restype = ctypes.c_int
    lib.hvd_ipc_open.argtypes = [ctypes.c_void_p, P(ctypes.c_void_p)]
    lib.hvd_ipc_close.argtypes = [ctypes.c_void_p]
    lib.hvd_cuda_error_string.restype = ctypes.c_char_p
    lib.hvd_cuda_error_string.argtypes = [ctypes.c_int]
    lib.hvd_memcpy_d2d.restype = ctypes.c_int
    lib.hvd_memcpy_d2d.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p]
    lib.hvd_pad_bytes.restype = ctypes.c_int
    lib.hvd_epoch_bytes.restype = ctypes.c_int
    lib._hvd_comm_declared = True


def _declare_symm(core):
    if getattr(core, "_hvd_symm_declared", False):
        return
    P = ctypes.POINTER
    U64 = ctypes.c_ulonglong
    core.hvd_symm_last_error.restype = ctypes.c_char_p
    core.hvd_symm_round_size.restype = ctypes.c_longlong
    core.hvd_symm_round_size.argtypes = [ctypes.c_int, ctypes.c_longlong, ctypes.c_int]
    core.hvd_symm_create.argtypes = [ctypes.c_int, ctypes.c_longlong, P(U64), P(ctypes.c_int)]
    core.hvd_symm_import.argtypes = [ctypes.c_int, P(U64)]
    core.hvd_symm_map.argtypes = [ctypes.c_int, U64, ctypes.c_longlong, P(U64)]
    core.hvd_symm_unmap.argtypes = [U64, ctypes.c_longlong]
    core.hvd_symm_release.argtypes = [U64]
    core.hvd_symm_mc_create.argtypes = [ctypes.c_int, ctypes.c_longlong, P(U64), P(ctypes.c_int)]
    core.hvd_symm_mc_add_device.argtypes = [U64, ctypes.c_int]
    core.hvd_symm_mc_bind.argtypes = [U64, U64, ctypes.c_longlong]
    core._hvd_symm_declared = True


class HvdBootMsg(ctypes.Structure):
    _fields_ = [("magic", ctypes.c_uint32), ("version", ctypes.c_uint16), ("kind", ctypes.c_uint16),
                ("src_rank", ctypes.c_int32), ("tag", ctypes.c_int32), ("payload", ctypes.c_uint64)]


BOOT_KIND_MEM, BOOT_KIND_MC = 1, 2


def _declare_boot(core):
    if getattr(core, "_hvd_boot_declared", False):
        return
    core.hvd_boot_open.restype = ctypes.c_void_p
    core.hvd_boot_open.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_double]
    core.hvd_boot_send.restype = ctypes.c_int
    core.hvd_boot_send.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_uint64, ctypes.c_int]
    core.hvd_boot_recv.restype = ctypes.c_int
    core.hvd_boot_recv.argtypes = [ctypes.c_void_p, ctypes.POINTER(HvdBootMsg), ctypes.POINTER(ctypes.c_int)]
    core.hvd_boot_close.argtypes = [ctypes.c_void_p]
    core.hvd_boot_last_error.restype = ctypes.c_char_p
    core._hvd_boot_declared = True


class _FdExchange:
    """All-to-all exchange of POSIX fds between the ranks of one host: the native bootstrap channel
    (``csrc/core/bootstrap.cpp``: fixed-layout ``HvdBootMsg`` + SCM_RIGHTS over unix datagram sockets)."""

    def __init__(self, rank: int, size: int, tag: str, timeout_s: float = 60.0):
        self.rank, self.size = rank, size
        self.core = native.core_lib()
        _declare_boot(self.core)
        self.dir = os.path.join(tempfile.gettempdir(), f"hvd-symm-{tag}")
        self.h = self.core.hvd_boot_open(self.dir.encode(), rank, size, float(timeout_s))
        if not self.h:
            raise RuntimeError(f"bootstrap channel: {self.core.hvd_boot_last_error().decode()}")

    def send(self, dst: int, fd: int, tag: int, kind: int = BOOT_KIND_MEM, payload: int = 0):
        if self.core.hvd_boot_send(self.h, dst, kind, tag, payload, fd) != 0:
            raise RuntimeError(f"bootstrap send to rank {dst}: {self.core.hvd_boot_last_error().decode()}")

    def recv(self):
        """-> (tag, source rank, fd)"""
        m, fd = HvdBootMsg(), ctypes.c_int(-1)
        if self.core.hvd_boot_recv(self.h, ctypes.byref(m), ctypes.byref(fd)) != 0:
            raise RuntimeError(f"bootstrap receive on rank {self.rank}: {self.core.hvd_boot_last_error().decode()}")
        self.last = m
        return int(m.tag), int(m.src_rank), int(fd.value)

    def close(self):
        if self.h:
            self.core.hvd_boot_close(self.h)
            self.h = None


class PeerContext:
    def __init__(self, rank: int, size: int, device: torch.device, emulated: bool = False,
                 backend: str = "auto", timeout_s: float = 60.0):
        if size > MAX_RANKS:
            raise RuntimeError(f"peer-memory collectives support up to {MAX_RANKS} ranks per node")
        self.rank, self.size, self.device = rank, size, device
        self.emulated = emulated
        self.lib = native.cuda_lib()
        _declare(self.lib)
        self.backend = "emulated" if emulated else backend
        self.multicast = False
        self._buffers: List[SymmBuffer] = []
        self._by_id = {}                     # native registry id -> SymmBuffer
        self._raw = []                       # (kind, handle/ptr, bytes) for teardown
        self._fdx: Optional[_FdExchange] = None
        self._alloc_id = 0
        self.timeout_ns = int(timeout_s * 1e9)
        self._err = torch.zeros(1, dtype=torch.int32).pin_memory() if device.type == "cuda" else None
        self._err_np = self._err.numpy() if self._err is not None else None      # cheap host-side polling
        if size == 1 and not emulated:
            self.backend = "local"           # single GPU: same kernels, nothing to exchange
        elif not emulated:
            self._pick_backend()
        pad_bytes = int(self.lib.hvd_pad_bytes())
        self.pad = self.alloc(pad_bytes, multicast=False)
        n_local = size if emulated else 1
        self._epochs = [torch.zeros(int(self.lib.hvd_epoch_bytes()) // 4, dtype=torch.int32, device=device)
                        for _ in range(n_local)]
        self.staging_bytes = int(os.environ.get("HVD_STAGING_BYTES", 8 << 20))
        self.staging = self.alloc(self.staging_bytes)
        self._ctx_cache = {}
        self._ll = None                      # receive buffers of the LL one-shot kernel (allocated on first use)

    # ------------------------------------------------------------------ construction
    @classmethod
    def emulate(cls, nranks: int, device=None) -> "PeerContext":
        device = device or torch.device("cuda", torch.cuda.current_device())
        return cls(0, nranks, torch.device(device), emulated=True)

    def _pick_backend(self):
        core = native.core_lib()
        _declare_symm(core)
        self.core = core
        want = self.backend
        dev = self.device.index
        vmm_ok = False
        if want in ("auto", "vmm") and core.hvd_symm_available():
            vmm_ok = core.hvd_symm_handle_type_supported(dev) == 1
        from .. import collectives
        oks = collectives.allgather_object(bool(vmm_ok))
        if all(oks) and want in ("auto", "vmm"):
            self.backend = "vmm"
            mc = core.hvd_symm_multicast_supported(dev) == 1 and os.environ.get("HVD_DISABLE_NVLS", "0") != "1"
            self.multicast = all(collectives.allgather_object(bool(mc)))
            tag = collectives.broadcast_object(f"{os.getpid()}-{os.environ.get('MASTER_PORT', '0')}", 0)
            self._fdx = _FdExchange(self.rank, self.size, tag)
            collectives.barrier()
        elif want == "vmm":
            raise RuntimeError("HVD_SYMM_BACKEND=vmm but the VMM/POSIX-fd path is not available on every rank")
        else:
            self.backend = "ipc"

    # ------------------------------------------------------------------ allocation
    def alloc(self, nbytes: int, multicast: Optional[bool] = None) -> SymmBuffer:
        """Collective: every rank must call with the same size, in the same order."""
        nbytes = (int(nbytes) + 255) // 256 * 256
        if self.emulated or self.backend == "local":
            buf = self._alloc_emulated(nbytes)
        elif self.backend == "vmm":
            buf = self._alloc_vmm(nbytes, self.multicast if multicast is None else (multicast and self.multicast))
        else:
            buf = self._alloc_ipc(nbytes)
        buf.tensor.zero_()
        if self.emulated:
            for r in range(self.size):
                buf.peer_tensor(r).zero_()
        torch.cuda.synchronize(self.device)
        if not self.emulated and self.size > 1:
            from .. import collectives
            collectives.barrier()
        self._buffers.append(buf)
        core = native.core_lib(required=False)
        if core is not None:
            # the native registry (csrc/core/runtime.cpp) knows where every rank's copy of this buffer is mapped
            arr = (ctypes.c_uint64 * len(buf.ptrs))(*buf.ptrs)
            buf.registry_id = core.hvd_registry_add(arr, len(buf.ptrs), buf.nbytes, buf.mc_ptr or 0, buf.local_index)
            self._by_id[buf.registry_id] = buf
        return buf

    def _alloc_emulated(self, nbytes: int) -> SymmBuffer:
        whole = torch.empty(nbytes * self.size // 4, dtype=torch.float32, device=self.device)
        base = whole.data_ptr()
        buf = SymmBuffer(nbytes, [base + r * nbytes for r in range(self.size)], 0, 0, self.device)
        buf._keep.append(whole)
        return buf

    def _chk_symm(self, rc: int, what: str):
        if rc != 0:
            raise RuntimeError(f"{what} failed: {self.core.hvd_symm_last_error().decode()}")

    def _alloc_vmm(self, nbytes: int, want_mc: bool) -> SymmBuffer:
        from .. import collectives
        core, dev = self.core, self.device.index
        U64 = ctypes.c_ulonglong
        size = int(core.hvd_symm_round_size(dev, nbytes, self.size if want_mc else 0))
        if size <= 0:
            raise RuntimeError("hvd_symm_round_size failed")
        h, fd = U64(), ctypes.c_int()
        self._chk_symm(core.hvd_symm_create(dev, size, ctypes.byref(h), ctypes.byref(fd)), "cuMemCreate/export")
        self._alloc_id += 1
        tag = self._alloc_id
        for r in range(self.size):
            if r != self.rank:
                self._fdx.send(r, fd.value, tag, kind=BOOT_KIND_MEM, payload=size)
        handles = {self.rank: h.value}
        for _ in range(self.size - 1):
            t, src, pfd = self._fdx.recv()
            if t != tag or self._fdx.last.kind != BOOT_KIND_MEM or self._fdx.last.payload != size:
                raise RuntimeError(f"symmetric alloc out of order: rank {src} sent tag {t} / {self._fdx.last.payload} B, "
                                   f"expected tag {tag} / {size} B (every rank must allocate in the same order)")
            ph = U64()
            self._chk_symm(core.hvd_symm_import(pfd, ctypes.byref(ph)), "cuMemImportFromShareableHandle")
            os.close(pfd)
            handles[src] = ph.value
        os.close(fd.value)
        ptrs = []
        for r in range(self.size):
            p = U64()
            self._chk_symm(core.hvd_symm_map(dev, handles[r], size, ctypes.byref(p)), "cuMemMap/SetAccess")
            ptrs.append(p.value)
            self._raw.append(("vmm", handles[r], p.value, size))
        mc_ptr = 0
        if want_mc:
            mc_ptr = self._bind_multicast(handles[self.rank], size, tag)
        collectives.barrier()
        return SymmBuffer(nbytes, ptrs, mc_ptr, self.rank, self.device)

    def _bind_multicast(self, mem_handle: int, size: int, tag: int) -> int:
        from .. import collectives
        core, dev = self.core, self.device.index
        U64 = ctypes.c_ulonglong
        mc = U64()
        ok = True
        mtag = tag | 0x40000000
        try:
            if self.rank == 0:
                fd = ctypes.c_int()
                rc = core.hvd_symm_mc_create(self.size, size, ctypes.byref(mc), ctypes.byref(fd))
                created = rc == 0
                created = collectives.broadcast_object(created, 0)
                if not created:
                    raise RuntimeError("cuMulticastCreate failed")
                for r in range(1, self.size):
                    self._fdx.send(r, fd.value, mtag, kind=BOOT_KIND_MC, payload=size)
                os.close(fd.value)
            else:
                if not collectives.broadcast_object(None, 0):
                    raise RuntimeError("cuMulticastCreate failed on rank 0")
                t, _src, pfd = self._fdx.recv()
                if t != mtag:
                    raise RuntimeError("multicast handle out of order")
                self._chk_symm(core.hvd_symm_import(pfd, ctypes.byref(mc)), "import multicast handle")
                os.close(pfd)
            self._chk_symm(core.hvd_symm_mc_add_device(mc.value, dev), "cuMulticastAddDevice")
        except RuntimeError as e:
            basics.log.warning("NVLS setup failed (%s); falling back to P2P variants", e)
            ok = False
        if not all(collectives.allgather_object(ok)):
            self.multicast = False
            return 0
        ok = core.hvd_symm_mc_bind(mc.value, mem_handle, size) == 0
        if not all(collectives.allgather_object(ok)):
            self.multicast = False
            return 0
        p = U64()
        ok = core.hvd_symm_map(dev, mc.value, size, ctypes.byref(p)) == 0
        if not all(collectives.allgather_object(ok)):
            self.multicast = False
            return 0
        self._raw.append(("mc", mc.value, p.value, size))
        return p.value

    def _alloc_ipc(self, nbytes: int) -> SymmBuffer:
        from .. import collectives
        p = ctypes.c_void_p()
        rc = self.lib.hvd_cuda_malloc(ctypes.byref(p), nbytes)
        if rc != 0:
            raise RuntimeError(f"cudaMalloc failed: {self.lib.hvd_cuda_error_string(rc).decode()}")
        h = ctypes.create_string_buffer(64)
        rc = self.lib.hvd_ipc_get_handle(p, h)
        if rc != 0:
            raise RuntimeError(f"cudaIpcGetMemHandle failed: {self.lib.hvd_cuda_error_string(rc).decode()}")
        handles = collectives.allgather_object(h.raw)
        ptrs = []
        for r in range(self.size):
            if r == self.rank:
                ptrs.append(p.value)
                continue
            q = ctypes.c_void_p()
            rc = self.lib.hvd_ipc_open(handles[r], ctypes.byref(q))
            if rc != 0:
                raise RuntimeError(f"cudaIpcOpenMemHandle(rank {r}) failed: "
                                   f"{self.lib.hvd_cuda_error_string(rc).decode()}")
            ptrs.append(q.value)
            self._raw.append(("ipc_peer", 0, q.value, nbytes))
        self._raw.append(("ipc_own", 0, p.value, nbytes))
        return SymmBuffer(nbytes, ptrs, 0, self.rank, self.device)

    def alloc_tensor(self, nbytes: int) -> torch.Tensor:
        return self.alloc(nbytes).tensor

    def buffer_of(self, t: torch.Tensor):
        """Return (SymmBuffer, element offset) when ``t`` lives inside a symmetric buffer."""
        p = t.data_ptr()
        core = native.core_lib(required=False)
        if core is not None and self._by_id:
            off = ctypes.c_uint64(0)
            b = self._by_id.get(core.hvd_registry_find(p, ctypes.byref(off)))
            if b is not None:
                return b, off.value // 4
            # (the registry is per process; an entry of another context may shadow nothing of ours - fall through)
        for b in self._buffers:
            base = b.ptrs[b.local_index]
            if base <= p < base + b.nbytes:
                return b, (p - base) // 4
        return None, 0

    # ------------------------------------------------------------------ launching
    def comm_ctx(self, buf: SymmBuffer) -> HvdCommCtx:
        key = id(buf)
        c = self._ctx_cache.get(key)
        if c is not None:
            return c
        c = HvdCommCtx()
        c.rank = -1 if self.emulated else self.rank
        c.nranks = self.size
        for r in range(self.size):
            c.base[r] = buf.ptrs[r]
            c.pad[r] = self.pad.ptrs[r]
            c.epoch[r] = self._epochs[r if self.emulated else 0].data_ptr()
        c.mc_base = buf.mc_ptr or None
        c.error = self._err.data_ptr()
        c.timeout_ns = self.timeout_ns
        self._ctx_cache[key] = c
        return c

    def _stream(self, stream=None) -> int:
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        return s.cuda_stream

    def launch_fused(self, buf: SymmBuffer, args: HvdFusedArgs, ctas: int, threads: int = 256, stream=None):
        if args.variant == NVLS and not buf.mc_ptr:
            raise RuntimeError("NVLS variant requested but this buffer has no multicast mapping")
        if self.emulated:
            ctas = max(1, min(ctas, 148 // self.size))
        rc = self.lib.hvd_fused_allreduce_launch(ctypes.byref(self.comm_ctx(buf)), ctypes.byref(args),
                                                 int(ctas), int(threads), self._stream(stream),
                                                 self.size if self.emulated else 0)
        if rc != 0:
            msg = self.lib.hvd_cuda_error_string(rc).decode() if rc > 0 else f"bad arguments ({rc})"
            raise RuntimeError(f"hvd_fused_allreduce_launch failed: {msg}")

    def launch_ll(self, buf: SymmBuffer, in_off: int, out_off: int, n: int, scale: float, stream=None):
        """Barrier-free LL one-shot allreduce of ``n`` (even) floats at ``in_off`` of ``buf`` into ``out_off``.
        Collective; the first call allocates the symmetric receive buffers (collective too)."""
        cap = (LL_MAX_BYTES // 4 + 1) // 2
        if self._ll is None:
            self._ll = self.alloc(2 * self.size * cap * 16, multicast=False)
        a = HvdLLArgs()
        a.in_off, a.out_off, a.n, a.scale, a.cap = in_off, out_off, n, scale, cap
        for r in range(self.size):
            a.ll[r] = self._ll.ptrs[r]
        rc = self.lib.hvd_ll_allreduce_launch(ctypes.byref(self.comm_ctx(buf)), ctypes.byref(a), self._stream(stream),
                                              self.size if self.emulated else 0)
        if rc != 0:
            msg = self.lib.hvd_cuda_error_string(rc).decode() if rc > 0 else f"bad arguments ({rc})"
            raise RuntimeError(f"hvd_ll_allreduce_launch failed: {msg}")

    def launch_bcast(self, buf: SymmBuffer, root: int, off: int, n: int, slot: int = SLOT_BCAST,
                     ctas: int = 0, stream=None):
        ctas = ctas or max(1, min(32, (n * 4 + 65535) // 65536))
        if self.emulated:
            ctas = max(1, min(ctas, 148 // self.size))
        rc = self.lib.hvd_bcast_launch(ctypes.byref(self.comm_ctx(buf)), slot, root, off, n, ctas, 512,
                                       self._stream(stream), self.size if self.emulated else 0)
        if rc != 0:
            raise RuntimeError(f"hvd_bcast_launch failed: {rc}")

    def pull(self, buf: SymmBuffer, src_rank: int, elem_off: int, n_elems: int, stream=None) -> None:
        """One-sided read: copy [elem_off, elem_off+n) of rank ``src_rank``'s copy of ``buf`` into this
        rank's copy (cudaMemcpyAsync on the peer-mapped address; torch never sees the peer pointer)."""
        dst = buf.ptrs[buf.local_index] + 4 * elem_off
        src = buf.ptrs[src_rank] + 4 * elem_off
        rc = self.lib.hvd_memcpy_d2d(dst, src, 4 * n_elems, self._stream(stream))
        if rc != 0:
            raise RuntimeError(f"peer pull failed: {self.lib.hvd_cuda_error_string(rc).decode()}")

    def check_error(self) -> None:
        """Raise if an in-kernel barrier timed out (names the peer that did not arrive)."""
        v = int(self._err_np[0])
        if v:
            self._err_np[0] = 0
            raise RuntimeError(f"peer-memory barrier timed out: rank {self.rank} waited for rank "
                               f"{(v & 0xff) - 1} on signal slot {v >> 8} (stalled or failed peer)")

    # ------------------------------------------------------------------ tensor-level collectives
    def _pick(self, nbytes: int):
        from . import variants
        cfg = basics.state().config
        forced = cfg.allreduce_variant if cfg else "auto"
        return variants.pick(nbytes, self.size, bool(self.staging.mc_ptr), forced,
                             cfg.comm_ctas if cfg else 0)

    def allreduce_(self, t: torch.Tensor, average: bool = True) -> torch.Tensor:
        """In-place allreduce of an arbitrary CUDA tensor through the staging buffer."""
        if t.dtype == torch.float64:
            # The kernels sum fp32.  A float64 tensor (the metric vector) travels as three fp32 pieces of
            # <= 18 significant bits each, so the fp32 sum of up to 8 ranks' pieces is exact whenever the
            # ranks' values are within a factor of 8 of each other (metrics are); recombined in fp64.
            pieces, rest = [], t.reshape(-1).clone()
            for _ in range(2):
                hi = (rest.to(torch.float32).view(torch.int32) & ~0x3F).view(torch.float32)
                pieces.append(hi)
                rest = rest - hi.to(torch.float64)
            pieces.append(rest.to(torch.float32))
            packed = torch.cat(pieces)
            self.allreduce_(packed, average=False)
            n = t.numel()
            tot = packed[:n].to(torch.float64) + packed[n:2 * n].to(torch.float64) + packed[2 * n:].to(torch.float64)
            if average:
                tot = tot / self.size
            t.copy_(tot.view(t.shape))
            return t
        buf, off = self.buffer_of(t)
        flat = t.reshape(-1) if t.is_contiguous() else None
        if buf is not None and flat is not None and t.dtype == torch.float32 and off % 4 == 0 \
                and t.numel() % 4 == 0:
            self._allreduce_range(buf, off, t.numel(), average)
            return t
        cap = self.staging_bytes // 8            # first half: inputs, second half: results
        src = t.reshape(-1).to(torch.float32) if (t.dtype != torch.float32 or not t.is_contiguous()) \
            else t.reshape(-1)
        n = src.numel()
        stage = self.staging.tensor
        out = src if src.data_ptr() == t.data_ptr() else torch.empty_like(src)
        for s in range(0, n, cap):
            m = min(cap, n - s)
            m4 = (m + 3) // 4 * 4
            stage[:m].copy_(src[s:s + m])
            if m4 > m:
                stage[m:m4].zero_()
            self._allreduce_range(self.staging, 0, m4, average, out_off=cap)
            out[s:s + m].copy_(stage[cap:cap + m])
        if out.data_ptr() != t.data_ptr():
            t.copy_(out.view(t.shape).to(t.dtype))
        return t

    def _allreduce_range(self, buf: SymmBuffer, off: int, n: int, average: bool, stream=None,
                         out_off: Optional[int] = None):
        variant, ctas = self._pick(n * 4)
        cfg = basics.state().config
        forced = cfg.allreduce_variant if cfg else "auto"
        if self.size > 1 and n * 4 <= LL_MAX_BYTES and n % 2 == 0 and off % 2 == 0 and \
                (forced == "ll" or (forced == "auto" and variant == ONESHOT)):
            self.launch_ll(buf, off, off if out_off is None else out_off, n, 1.0 / self.size if average else 1.0,
                           stream=stream)
            return
        if variant == NVLS and not buf.mc_ptr:
            variant = TWOSHOT
        if out_off is None and variant == ONESHOT:
            # in place: a one-shot rank would overwrite data its peers are still reading;
            # two-shot is safe in place (each element is read only by its single owner CTA)
            variant = TWOSHOT
        a = HvdFusedArgs()
        a.start, a.end = 0, n
        a.grad_off = off
        a.out_off = off if out_off is None else out_off
        a.param_off = a.slot0_off = a.slot1_off = 0
        a.hyper = None
        a.scale = 1.0 / self.size if average else 1.0
        a.kind, a.variant, a.slot, a.exit_barrier = 0, variant, SLOT_STAGING, 1
        self.launch_fused(buf, a, ctas, stream=stream)

    def broadcast_(self, t: torch.Tensor, root: int) -> torch.Tensor:
        buf, off = self.buffer_of(t)
        if buf is not None and t.is_contiguous() and t.dtype == torch.float32 and off % 4 == 0 \
                and t.numel() % 4 == 0:
            self.launch_bcast(buf, root, off, t.numel())
            return t
        cap = self.staging_bytes // 4
        raw = t.contiguous().view(torch.uint8).reshape(-1)
        nb = raw.numel()
        stage8 = self.staging.tensor.view(torch.uint8)
        out = raw if raw.data_ptr() == t.data_ptr() else torch.empty_like(raw)
        for s in range(0, nb, cap * 4):
            m = min(cap * 4, nb - s)
            m16 = (m + 15) // 16 * 16
            if self.rank == root:
                stage8[:m].copy_(raw[s:s + m])
            self.launch_bcast(self.staging, root, 0, m16 // 4, slot=SLOT_BCAST)
            out[s:s + m].copy_(stage8[:m])
        if out.data_ptr() != t.data_ptr():
            t.copy_(out.view(t.dtype).view(t.shape))
        return t

    def close(self):
        if getattr(self, "_closed", False):
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
        except Exception:
            pass
        self._buffers.clear()
        core = native.core_lib(required=False)
        if core is not None:
            for bid in self._by_id:
                core.hvd_registry_remove(bid)
        self._by_id.clear()
        if self._fdx is not None:
            self._fdx.close()
        # VMM/IPC mappings are released with the process; explicit unmap order across ranks is
        # not needed for correctness and risks tearing down memory a peer kernel still reads.


def available() -> bool:
    """True when this process can use the peer-memory kernels (CUDA, size>1, one node)."""
    global _unavailable_reason
    st = basics.state()
    if not st.initialized or st.device is None or st.device.type != "cuda":
        return False
    if st.config.allreduce_variant == "nccl":
        return False
    if st.size > MAX_RANKS or st.local_size != st.size:
        _unavailable_reason = "peer memory needs all ranks on one NVSwitch node"
        return False
    if _ctx is not None:
        return True
    return _unavailable_reason is None


def context() -> PeerContext:
    """The process-wide context (created collectively on first use)."""
    global _ctx, _unavailable_reason
    if _ctx is None:
        st = basics.state()
        try:
            _ctx = PeerContext(st.rank, st.size, st.device, backend=st.config.symm_backend,
                               timeout_s=st.config.barrier_abort_time())
        except Exception as e:
            _unavailable_reason = str(e)
            raise
        st.peer = _ctx
    return _ctx


def reset():
    global _ctx, _unavailable_reason
    _ctx = None
    _unavailable_reason = None
