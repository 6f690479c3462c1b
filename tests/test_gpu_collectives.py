"""GPU tests of the peer-memory collectives.

Single-GPU: N virtual ranks emulated in ONE cooperative launch (blockIdx.y = rank) - separate
waiting launches on one GPU are unsafe (B200_PROFILING.md).  Multi-GPU: real processes, one
per GPU, when >= 2 GPUs are visible.
"""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _need_gpu():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device is visible")


@pytest.fixture(scope="module")
def ctx4():
    _need_gpu()
    from horovod_distributed_example_b200.parallel.peer import PeerContext
    return PeerContext.emulate(4)


@pytest.fixture(scope="module")
def ctx8():
    _need_gpu()
    from horovod_distributed_example_b200.parallel.peer import PeerContext
    return PeerContext.emulate(8)


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("n", [4, 1000, 262144 + 12])
def test_emulated_allreduce_matches_sum(ctx4, variant, n):
    from horovod_distributed_example_b200.parallel import peer as P
    ctx = ctx4
    n4 = (n + 3) // 4 * 4
    buf = ctx.staging
    torch.manual_seed(n)
    vals = torch.randn(ctx.size, n4, device="cuda")
    for r in range(ctx.size):
        buf.peer_tensor(r)[:n4].copy_(vals[r])
    a = P.HvdFusedArgs()
    a.start, a.end = 0, n4
    a.grad_off, a.out_off = 0, (0 if variant else 1 << 20)   # one-shot must not run in place
    a.scale, a.kind, a.variant, a.slot, a.exit_barrier = 1.0 / ctx.size, 0, variant, P.SLOT_STAGING, 1
    for _ in range(3):                      # repeated launches exercise the monotonic epochs
        for r in range(ctx.size):
            buf.peer_tensor(r)[:n4].copy_(vals[r])
        ctx.launch_fused(buf, a, ctas=8)
    torch.cuda.synchronize()
    ctx.check_error()
    ref = vals[0].clone()
    for r in range(1, ctx.size):
        ref += vals[r]                       # same fixed order as the kernel
    ref *= 1.0 / ctx.size
    oo = a.out_off
    outs = [buf.peer_tensor(r)[oo:oo + n4].clone() for r in range(ctx.size)]
    for o in outs:
        assert torch.equal(o, outs[0])       # bit-identical on every rank
    torch.testing.assert_close(outs[0], ref, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("n", [2, 1000, 32768])
def test_emulated_ll_allreduce_matches_sum(ctx4, n):
    """The barrier-free LL one-shot kernel (small messages): in place and out of place, repeated launches (epoch
    parity double-buffering), bit-identical on every rank."""
    ctx = ctx4
    buf = ctx.staging
    torch.manual_seed(n)
    vals = torch.randn(ctx.size, n, device="cuda")
    ref = vals[0].clone()
    for r in range(1, ctx.size):
        ref += vals[r]
    ref *= 1.0 / ctx.size
    for rep in range(5):
        out_off = 0 if rep % 2 else (1 << 18)
        for r in range(ctx.size):
            buf.peer_tensor(r)[:n].copy_(vals[r])
        ctx.launch_ll(buf, 0, out_off, n, 1.0 / ctx.size)
        torch.cuda.synchronize()
        ctx.check_error()
        outs = [buf.peer_tensor(r)[out_off:out_off + n].clone() for r in range(ctx.size)]
        for o in outs:
            assert torch.equal(o, outs[0])
        torch.testing.assert_close(outs[0], ref, rtol=1e-6, atol=1e-6)


def test_emulated_broadcast(ctx4):
    ctx = ctx4
    buf = ctx.staging
    n = 4096
    for r in range(ctx.size):
        buf.peer_tensor(r)[:n].fill_(float(r + 1))
    ctx.launch_bcast(buf, 2, 0, n)
    torch.cuda.synchronize()
    for r in range(ctx.size):
        assert torch.all(buf.peer_tensor(r)[:n] == 3.0)


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("opt_name", ["adadelta", "adam", "sgd"])
def test_emulated_fused_update_matches_cpu_oracle(ctx8, variant, opt_name):
    """allreduce + 1/N + optimizer update in one kernel == CPU oracle on the mean gradient;
    every replica's params bit-identical; two-shot state sharded by owner."""
    from horovod_distributed_example_b200.keras import optimizers as O
    from horovod_distributed_example_b200.parallel import buckets as B
    from horovod_distributed_example_b200.parallel.arena import ParamArena
    from horovod_distributed_example_b200.parallel.fused import FusedReducer
    ctx = ctx8
    N = ctx.size
    specs = [("d2/b", (10,)), ("d2/k", (128, 10)), ("d1/b", (128,)), ("d1/k", (300, 128)), ("c/k", (3, 3, 4, 8))]
    mk = {"adadelta": lambda: O.Adadelta(2.0), "adam": lambda: O.Adam(0.01), "sgd": lambda: O.SGD(0.1, momentum=0.9)}
    opt = mk[opt_name]()
    arena = ParamArena(specs, "cuda", n_slots=opt.n_slots, allocator=ctx.alloc_tensor)
    opt.bind(arena)
    buckets = B.plan(arena, 64 << 20, ["d1/k"])
    assert len(buckets) == 2
    red = FusedReducer(arena, opt, buckets, ctx, variants=[variant, variant])
    buf, base = red.buf, red.base_off
    n = arena.n
    cpu_opt = mk[opt_name]()
    cpu = ParamArena(specs, "cpu", n_slots=opt.n_slots)
    cpu_opt.bind(cpu)
    g = torch.Generator().manual_seed(5)
    p0 = torch.randn(n, generator=g)
    for s in arena.specs:                      # keep padding zero
        pass
    cpu.params.copy_(p0)
    for r in range(N):
        buf.peer_tensor(r)[base:base + n].copy_(p0)
    for step in range(3):
        grads = torch.randn(N, n, generator=g)
        for r in range(N):
            buf.peer_tensor(r)[base + arena.grad_offset(0): base + arena.grad_offset(0) + n].copy_(grads[r])
        opt.iterations += 1
        red.run_all()
        opt._push_hyper()
        torch.cuda.synchronize()
        ctx.check_error()
        mean = grads[0].clone()
        for r in range(1, N):
            mean += grads[r]
        cpu.grads.copy_(mean * (1.0 / N))
        cpu_opt.apply_gradients(cpu)
    reps = [buf.peer_tensor(r)[base:base + n].cpu() for r in range(N)]
    for rp in reps:
        assert torch.equal(rp, reps[0])
    torch.testing.assert_close(reps[0], cpu.params, rtol=2e-5, atol=2e-6)
    red.materialize_state()                    # rank 0 pulls the shards it does not own
    for k in range(opt.n_slots):
        got = arena.slots[k].cpu()
        torch.testing.assert_close(got, cpu.slots[k], rtol=2e-5, atol=1e-7)


def test_barrier_timeout_sets_error_word():
    """Fault injection: a 2-rank context where only 1 rank's blocks run must not hang - the
    bounded spin raises the error word naming the missing peer (stall inspector)."""
    _need_gpu()
    from horovod_distributed_example_b200.parallel import peer as P
    ctx = P.PeerContext.emulate(2)
    ctx.timeout_ns = int(0.2e9)
    ctx._ctx_cache.clear()
    a = P.HvdFusedArgs()
    a.start, a.end, a.grad_off, a.out_off = 0, 1024, 0, 0
    a.scale, a.kind, a.variant, a.slot, a.exit_barrier = 0.5, 0, 0, 5, 0
    c = ctx.comm_ctx(ctx.staging)
    import ctypes
    c2 = P.HvdCommCtx.from_buffer_copy(c)
    c2.rank = 0                                     # production mode: rank 1 never launches
    rc = ctx.lib.hvd_fused_allreduce_launch(ctypes.byref(c2), ctypes.byref(a), 1, 256,
                                            torch.cuda.current_stream().cuda_stream, 0)
    assert rc == 0
    torch.cuda.synchronize()
    with pytest.raises(RuntimeError, match="waited for rank 1"):
        ctx.check_error()


def _run_multi(nproc, body, timeout=300, extra_env=None):
    env = dict(os.environ)
    env["HVD_TEST_BODY"] = body
    env.update(extra_env or {})
    return subprocess.run([sys.executable, "-m", "horovod_distributed_example_b200.run", "-np", str(nproc),
                           "--timeout", str(timeout), os.path.join(ROOT, "tests", "_gpu_workers.py")],
                          cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout + 60)


@pytest.mark.multigpu
@pytest.mark.parametrize("body", ["collectives", "train_replicas", "torchmodel_overlap", "torchmodel_bn_vs_ddp"])
def test_multi_gpu(body):
    _need_gpu()
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _run_multi(min(n, 8), body)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("OK") >= min(n, 8)


@pytest.mark.multigpu
def test_fault_injection_on_real_ranks():
    """SURVEY section 5.3 end to end: rank 1 hangs at step 3 (HVD_FAULT_INJECT=1:3).  The other ranks' in-kernel
    waits give up after the shutdown time, the host error names rank 1, and hvdrun kills the group."""
    _need_gpu()
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    import time
    t0 = time.time()
    r = _run_multi(min(n, 8), "fault_inject", timeout=120,
                   extra_env={"HVD_FAULT_INJECT": "1:3", "HVD_STALL_CHECK_TIME_SECONDS": "1",
                              "HVD_STALL_SHUTDOWN_TIME_SECONDS": "3"})
    out = r.stdout + r.stderr
    assert r.returncode != 0, out[-3000:]
    # (a step's loss is published before its tail kernel runs, so step 3 itself may still report; step 4 cannot)
    assert "rank 0: step 2 done" in out and "rank 0: step 4 done" not in out, out[-3000:]
    assert "FAULT-DETECTED rank 0" in out and "waited for rank 1" in out, out[-3000:]
    assert "terminating the job" in out, out[-3000:]
    assert "UNEXPECTED" not in out
    assert time.time() - t0 < 100
