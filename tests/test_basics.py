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
