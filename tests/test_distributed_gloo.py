"""BASELINE.json config 2: world_size=2 on CPU/gloo - broadcast + metric-average plumbing."""
import pytest

from conftest import run_workers


def _ok(outs):
    for rc, out in outs:
        assert rc == 0 and out.strip().endswith("OK"), out


def test_collectives_ws2():
    _ok(run_workers("collectives"))


def test_collectives_ws3():
    _ok(run_workers("collectives", world_size=3))


def test_broadcast_and_average_gradients_ws2():
    _ok(run_workers("broadcast_and_grads"))


def test_tf2_script_flow_ws2(tmp_path):
    _ok(run_workers("tf2_script_flow", env={"HVD_TEST_OUT": str(tmp_path)}))
    assert (tmp_path / "tb" / "train").exists()


def test_bucket_plan_mismatch_detected():
    _ok(run_workers("plan_mismatch"))


def test_resume_from_rank0_checkpoint_ws2(tmp_path):
    _ok(run_workers("resume_from_checkpoint", env={"HVD_TEST_OUT": str(tmp_path)}))


def test_hvd_load_model_and_broadcast_optimizer_state_ws2(tmp_path):
    _ok(run_workers("hvd_load_model_and_optimizer_state", env={"HVD_TEST_OUT": str(tmp_path)}))
