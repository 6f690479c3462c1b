"""Build and load the native libraries (in-tree, C ABI, bound with ctypes).

* ``lib/libhvd_core.so``  - host C++ runtime (HDF5, event files, timeline, bucket planner,
  symmetric-memory arena).  Built with g++; links the CUDA driver API lazily via dlopen.
* ``lib/libhvd_cuda.so``  - sm_100a kernels (collectives + model ops).  Built with
  ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo``.

The libraries are built in-tree so they travel to the GPU box with the snapshot; nothing
is JIT-compiled at import.  On a GPU box a missing CUDA library is a hard error: the
CUDA path must never silently fall back to eager PyTorch.
"""
from __future__ import annotations

import ctypes
import glob
import os
import shutil
import subprocess
import sys
import threading
from typing import List, Optional

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_DIR = os.path.join(PKG_DIR, "lib")
INCLUDE_DIR = os.path.join(os.path.dirname(PKG_DIR), "include")
CORE_SO = os.path.join(LIB_DIR, "libhvd_core.so")
CUDA_SO = os.path.join(LIB_DIR, "libhvd_cuda.so")

CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
NVCC = os.path.join(CUDA_HOME, "bin", "nvcc")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]

_lock = threading.Lock()
_core = None
_cuda = None


def _sources(sub: str, exts) -> List[str]:
    out = []
    for e in exts:
        out += glob.glob(os.path.join(CSRC, sub, "*" + e))
    return sorted(out)


def _newer(target: str, deps: List[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd: List[str], verbose: bool):
    if verbose:
        print("+", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"build failed: {' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
    if verbose and (r.stdout or r.stderr):
        print(r.stdout, r.stderr, flush=True)
    return r


def build_core(force: bool = False, verbose: bool = False) -> str:
    srcs = _sources("core", (".cpp",)) + _sources("io", (".cpp",))
    hdrs = glob.glob(os.path.join(INCLUDE_DIR, "*.h")) + glob.glob(os.path.join(CSRC, "*", "*.h"))
    os.makedirs(LIB_DIR, exist_ok=True)
    if force or _newer(CORE_SO, srcs + hdrs):
        cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wall", "-I", INCLUDE_DIR,
               "-I", os.path.join(CUDA_HOME, "include"), "-o", CORE_SO] + srcs + ["-ldl", "-lpthread"]
        _run(cmd, verbose)
    return CORE_SO


def build_cuda(force: bool = False, verbose: bool = False, ptxas_info: bool = False) -> str:
    srcs = _sources("comm", (".cu",)) + _sources("kernels", (".cu",))
    hdrs = glob.glob(os.path.join(INCLUDE_DIR, "*.h")) + glob.glob(os.path.join(CSRC, "*", "*.cuh")) \
        + glob.glob(os.path.join(CSRC, "*", "*.h"))
    os.makedirs(LIB_DIR, exist_ok=True)
    if not os.path.exists(NVCC):
        raise RuntimeError(f"nvcc not found at {NVCC}")
    objs = []
    objdir = os.path.join(PKG_DIR, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for s in srcs:
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        objs.append(o)
        if force or _newer(o, [s] + hdrs):
            cmd = [NVCC, "-O3", "-std=c++17", "-lineinfo"]
            cmd += ARCH_FLAGS + ["-Xcompiler", "-fPIC", "-I", INCLUDE_DIR, "-I", os.path.join(CSRC),
                                 "-c", s, "-o", o]
            if ptxas_info:
                cmd += ["-Xptxas", "-v"]
            if verbose:
                print("+", " ".join(cmd), flush=True)
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)))
    for cmd, p in procs:
        out, err = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"build failed: {' '.join(cmd)}\n{out}\n{err}")
        if (verbose or ptxas_info) and (out or err):
            print(out, err, flush=True)
    if force or procs or _newer(CUDA_SO, objs):
        cmd = [NVCC, "-shared", "-o", CUDA_SO] + objs + ARCH_FLAGS + ["-lcudart"]
        _run(cmd, verbose)
    return CUDA_SO


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_core(force, verbose)
    build_cuda(force, verbose)


def core_lib(required: bool = True):
    """Load ``libhvd_core.so`` (building it if the toolchain is there and it is stale)."""
    global _core
    with _lock:
        if _core is not None:
            return _core
        try:
            if shutil.which("g++"):
                build_core()
            _core = ctypes.CDLL(CORE_SO, mode=ctypes.RTLD_GLOBAL)
            _declare_core(_core)
        except Exception as e:
            if required:
                raise RuntimeError(f"libhvd_core.so unavailable: {e}") from e
            return None
        return _core


def cuda_lib(required: bool = True):
    """Load ``libhvd_cuda.so``.  Never builds implicitly on a GPU box (the .so ships with
    the snapshot); raises loudly when it is missing and ``required``."""
    global _cuda
    with _lock:
        if _cuda is not None:
            return _cuda
        try:
            if not os.path.exists(CUDA_SO):
                build_cuda()
            _cuda = ctypes.CDLL(CUDA_SO, mode=ctypes.RTLD_GLOBAL)
            _declare_cuda(_cuda)
        except Exception as e:
            if required:
                raise RuntimeError(
                    f"libhvd_cuda.so unavailable ({e}); run `python -c 'import __graft_entry__ as g; "
                    "g.build()'` - the CUDA path never falls back to eager PyTorch") from e
            return None
        return _cuda


c_i64, c_i32, c_vp, c_cp = ctypes.c_int64, ctypes.c_int, ctypes.c_void_p, ctypes.c_char_p
P_I64 = ctypes.POINTER(ctypes.c_int64)
P_I32 = ctypes.POINTER(ctypes.c_int)


def _sig(lib, name, restype, argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = argtypes


def _declare_core(lib):
    _sig(lib, "hvd_plan_buckets", c_i64, [P_I64, P_I64, c_i64, c_i64, c_i64, P_I64, c_i64, P_I64])
    _sig(lib, "hvd_fnv1a", ctypes.c_uint64, [ctypes.c_uint64, c_cp, c_i64])
    _sig(lib, "hvd_crc32c", ctypes.c_uint32, [c_cp, c_i64])
    _sig(lib, "hvd_ev_open", c_vp, [c_cp])
    _sig(lib, "hvd_ev_scalar", c_i32, [c_vp, c_cp, ctypes.c_double, c_i64, ctypes.c_double])
    _sig(lib, "hvd_ev_flush", None, [c_vp])
    _sig(lib, "hvd_ev_close", None, [c_vp])
    _sig(lib, "hvd_tl_open", c_vp, [c_cp])
    _sig(lib, "hvd_tl_event", c_i32, [c_vp, c_cp, c_cp, ctypes.c_char, c_i64, c_i64])
    _sig(lib, "hvd_tl_now_us", c_i64, [c_vp])
    _sig(lib, "hvd_tl_close", None, [c_vp])
    _sig(lib, "hvd_log_set_level", None, [c_i32])
    _sig(lib, "hvd_log_level", c_i32, [])
    _sig(lib, "hvd_log_set_rank", None, [c_i32])
    _sig(lib, "hvd_log_parse_level", c_i32, [c_cp])
    _sig(lib, "hvd_log", None, [c_i32, c_cp])
    _sig(lib, "hvd_nvtx_push", c_i32, [c_cp])
    _sig(lib, "hvd_nvtx_pop", c_i32, [])
    _sig(lib, "hvd_nvtx_mark", None, [c_cp])
    _sig(lib, "hvd_core_discover_env", c_i32, [P_I32, ctypes.POINTER(c_cp)])
    _sig(lib, "hvd_core_init", c_i32, [])
    _sig(lib, "hvd_core_is_initialized", c_i32, [])
    _sig(lib, "hvd_core_rank", c_i32, [])
    _sig(lib, "hvd_core_size", c_i32, [])
    _sig(lib, "hvd_core_local_rank", c_i32, [])
    _sig(lib, "hvd_core_local_size", c_i32, [])
    _sig(lib, "hvd_core_launcher", c_cp, [])
    _sig(lib, "hvd_core_shutdown", None, [])
    _sig(lib, "hvd_registry_add", c_i32, [ctypes.POINTER(ctypes.c_uint64), c_i32, ctypes.c_uint64, ctypes.c_uint64, c_i32])
    _sig(lib, "hvd_registry_find", c_i32, [ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)])
    _sig(lib, "hvd_registry_peer_ptr", ctypes.c_uint64, [c_i32, c_i32, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)])
    _sig(lib, "hvd_registry_remove", c_i32, [c_i32])
    _sig(lib, "hvd_registry_count", c_i32, [])
    _sig(lib, "hvd_registry_clear", None, [])
    _sig(lib, "hvd_h5_last_error", c_cp, [])
    _sig(lib, "hvd_h5w_create", c_vp, [])
    _sig(lib, "hvd_h5w_free", None, [c_vp])
    _sig(lib, "hvd_h5w_group", c_i32, [c_vp, c_cp])
    _sig(lib, "hvd_h5w_dataset", c_i32, [c_vp, c_cp, c_i32, c_i32, P_I64, c_vp])
    _sig(lib, "hvd_h5w_attr_str", c_i32, [c_vp, c_cp, c_cp, c_cp, c_i64])
    _sig(lib, "hvd_h5w_attr_strs", c_i32, [c_vp, c_cp, c_cp, c_cp, c_i64, c_i64])
    _sig(lib, "hvd_h5w_attr_num", c_i32, [c_vp, c_cp, c_cp, c_i32, c_i32, P_I64, c_vp])
    _sig(lib, "hvd_h5w_write", c_i32, [c_vp, c_cp])
    _sig(lib, "hvd_h5r_open", c_vp, [c_cp])
    _sig(lib, "hvd_h5r_close", None, [c_vp])
    _sig(lib, "hvd_h5r_kind", c_i32, [c_vp, c_cp])
    _sig(lib, "hvd_h5r_list", c_i64, [c_vp, c_cp, c_cp, c_i64])
    _sig(lib, "hvd_h5r_dataset_info", c_i32, [c_vp, c_cp, P_I32, P_I32, P_I64, P_I64])
    _sig(lib, "hvd_h5r_dataset_read", c_i32, [c_vp, c_cp, c_vp, c_i64])
    _sig(lib, "hvd_h5r_attr_list", c_i64, [c_vp, c_cp, c_cp, c_i64])
    _sig(lib, "hvd_h5r_attr_info", c_i32, [c_vp, c_cp, c_cp, P_I32, P_I32, P_I64, P_I64])
    _sig(lib, "hvd_h5r_attr_read", c_i32, [c_vp, c_cp, c_cp, c_vp, c_i64])


def _declare_cuda(lib):
    # signatures are attached next to their Python wrappers (ops/*.py, parallel/fused.py)
    pass


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
    print("built", CORE_SO, CUDA_SO)
