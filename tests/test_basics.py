import os

import pytest

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import basics
from horovod_distributed_example_b200.config import Config, model_paths


def test_uninitialised_raises_valueerror():
    hvd.shutdown()
    for fn in (hvd.size, hvd.rank, hvd.local_rank):
        with pytest.raises(ValueError):
            fn()


def test_singleton_init_no_launcher(hvd_single):
    assert (hvd_single.size(), hvd_single.rank(), hvd_single.local_rank()) == (1, 0, 0)
    hvd_single.init()          # idempotent
    assert hvd_single.size() == 1
    assert hvd_single.local_size() == 1 and hvd_single.cross_size() == 1


@pytest.mark.parametrize("env,expect", [
    ({"HVD_RANK": "3", "HVD_SIZE": "8", "HVD_LOCAL_RANK": "3"}, (3, 8, 3, 8)),
    ({"OMPI_COMM_WORLD_RANK": "1", "OMPI_COMM_WORLD_SIZE": "4", "OMPI_COMM_WORLD_LOCAL_RANK": "1",
      "OMPI_COMM_WORLD_LOCAL_SIZE": "2"}, (1, 4, 1, 2)),
    ({"RANK": "5", "WORLD_SIZE": "8", "LOCAL_RANK": "5", "LOCAL_WORLD_SIZE": "8"}, (5, 8, 5, 8)),
    ({}, (0, 1, 0, 1)),
])
def test_env_contract(env, expect):
    assert basics.discover_env(env) == expect


def test_env_priority_hvd_over_torchrun():
    assert basics.discover_env({"HVD_RANK": "0", "HVD_SIZE": "2", "RANK": "1", "WORLD_SIZE": "4"})[:2] == (0, 2)


def test_bad_rank_rejected():
    with pytest.raises(ValueError):
        basics.discover_env({"HVD_RANK": "4", "HVD_SIZE": "4"})


def test_config_aliases(monkeypatch):
    monkeypatch.setenv("HOROVOD_FUSION_THRESHOLD", "1048576")
    monkeypatch.setenv("HVD_ALLREDUCE_VARIANT", "twoshot")
    monkeypatch.setenv("HVD_FAULT_INJECT", "1:7")
    c = Config.from_env()
    assert c.fusion_threshold == 1 << 20 and c.allreduce_variant == "twoshot"
    assert c.parsed_fault() == (1, 7)
    monkeypatch.setenv("HVD_ALLREDUCE_VARIANT", "ring")
    with pytest.raises(ValueError):
        Config.from_env()


def test_model_paths(monkeypatch, tmp_path):
    monkeypatch.setenv("PS_MODEL_PATH", str(tmp_path))
    md, ed = model_paths()
    assert md == str(tmp_path / "horovod-mnist") and ed == str(tmp_path)


def test_size1_collectives_are_identity(hvd_single):
    import torch
    t = torch.arange(4.0)
    assert torch.equal(hvd_single.allreduce(t), t)
    assert torch.equal(hvd_single.broadcast(t, 0), t)
    assert hvd_single.allreduce(3.0) == 3.0


def test_native_runtime_core_state_and_registry(monkeypatch):
    """csrc/core/runtime.cpp: the C API behind hvd.init()/rank()/size() (Horovod's horovod_init/horovod_rank/...)
    reads the launcher contracts itself, and the registry maps a local address back to its symmetric buffer."""
    import ctypes
    from horovod_distributed_example_b200.ops import native
    core = native.core_lib()
    core.hvd_core_shutdown()
    assert core.hvd_core_rank() == -1 and core.hvd_core_size() == -1            # before init, like Horovod's C API
    for k in ("HVD_RANK", "HVD_SIZE", "RANK", "WORLD_SIZE", "PMI_RANK", "PMI_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "3")
    monkeypatch.setenv("OMPI_COMM_WORLD_SIZE", "8")
    monkeypatch.setenv("OMPI_COMM_WORLD_LOCAL_RANK", "3")
    monkeypatch.setenv("OMPI_COMM_WORLD_LOCAL_SIZE", "8")
    assert core.hvd_core_init() == 0 and core.hvd_core_init() == 0             # idempotent
    assert (core.hvd_core_rank(), core.hvd_core_size(), core.hvd_core_local_rank()) == (3, 8, 3)
    assert core.hvd_core_launcher() == b"openmpi"
    monkeypatch.setenv("HVD_RANK", "9")
    monkeypatch.setenv("HVD_SIZE", "4")
    v = (ctypes.c_int * 4)()
    assert core.hvd_core_discover_env(v, None) == -1                            # rank outside [0, size)
    core.hvd_core_shutdown()
    ptrs = (ctypes.c_uint64 * 2)(0x10000, 0x90000)
    bid = core.hvd_registry_add(ptrs, 2, 0x4000, 0, 1)
    off = ctypes.c_uint64(0)
    assert core.hvd_registry_find(0x90100, ctypes.byref(off)) == bid and off.value == 0x100
    assert core.hvd_registry_find(0x10100, ctypes.byref(off)) == 0              # the peer's copy is not a local address
    assert core.hvd_registry_peer_ptr(bid, 0, None, None) == 0x10000
    core.hvd_registry_clear()
    assert core.hvd_registry_count() == 0
