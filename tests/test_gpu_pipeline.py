"""GPU tests of the pipelined data-parallel step (csrc/comm/pipeline_tail.cu + tc_gemm EPI_RS_PUSH).

Single GPU: N virtual ranks.  The producing GEMM is launched once per virtual rank (those launches do not wait
on one another), the tail kernel - whose CTAs wait for other ranks - runs as ONE cooperative launch with
blockIdx.y = rank.  Oracle: the mean gradient in fp32 on the CPU + the Keras-math optimizers.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _need_gpu():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device is visible")


ROWS = 1536          # 12 tiles of 128 rows: divisible by 1, 2, 3, 4, 6 (and 12)
SPECS = [("d2/b", (10,)), ("d2/k", (128, 10)), ("d1/b", (128,)), ("d1/k", (ROWS, 128)), ("c2/b", (64,)),
         ("c2/k", (3, 3, 32, 64)), ("c1/b", (32,)), ("c1/k", (3, 3, 1, 32))]


@pytest.mark.parametrize("N,flag_kernel", [(1, 1), (2, 1), (3, 1), (4, 1), (2, 0)])
@pytest.mark.parametrize("opt_name", ["adadelta", "adam"])
def test_emulated_pipelined_step_matches_cpu_oracle(N, flag_kernel, opt_name, monkeypatch):
    _need_gpu()
    monkeypatch.setenv("HVD_RS_FLAG_KERNEL", str(flag_kernel))
    from horovod_distributed_example_b200.keras import optimizers as O
    from horovod_distributed_example_b200.ops.gemm import EPI_RS_PUSH, tc_gemm
    from horovod_distributed_example_b200.parallel.arena import ParamArena
    from horovod_distributed_example_b200.parallel.peer import PeerContext
    from horovod_distributed_example_b200.parallel.pipeline import PipelinedStep
    ctx = PeerContext.emulate(N)
    mk = {"adadelta": lambda: O.Adadelta(2.0), "adam": lambda: O.Adam(0.01)}
    opt = mk[opt_name]()
    arena = ParamArena(SPECS, "cuda", n_slots=opt.n_slots, double_grads=True, allocator=ctx.alloc_tensor)
    opt.bind(arena)
    buf, base = ctx.buffer_of(arena.buffer)
    big = arena.spec("d1/k")
    w0, w1 = big.offset, big.offset + big.numel
    pipe = PipelinedStep(ctx, buf, base, arena, opt.kind, (0, w0), (w0, w1), (w1, arena.n))
    n = arena.n
    cpu_opt = mk[opt_name]()
    cpu = ParamArena(SPECS, "cpu", n_slots=opt.n_slots)
    cpu_opt.bind(cpu)
    g = torch.Generator().manual_seed(7)
    p0 = torch.randn(n, generator=g)
    cpu.params.copy_(p0)
    for r in range(N):
        buf.peer_tensor(r)[base:base + n].copy_(p0)
    B = 128
    for step in range(4):
        which = step & 1
        go = base + arena.grad_offset(which)
        grads = torch.randn(N, n, generator=g) * 0.1
        # the big gradient is produced by the GEMM  X^T[ROWS,B] * dY[B,128]  of every rank and pushed to its owner
        # (operands rounded to TF32-representable values: the tensor-core product then matches fp64 up to the
        # fp32 accumulation order, so the comparison can stay tight even for Adam's sign-like update)
        tf32 = lambda t: (t.view(torch.int32) & ~0x1FFF).view(torch.float32)
        xs = tf32(torch.randn(N, B, ROWS, generator=g)).cuda()
        dys = tf32(torch.randn(N, B, 128, generator=g) * 0.1).cuda()
        for r in range(N):
            buf.peer_tensor(r)[go: go + n].copy_(grads[r])
            tc_gemm(xs[r], dys[r], ROWS, 128, B, a_mn=True, b_mn=True, bn=128, epilogue=EPI_RS_PUSH, out=None,
                    rs=pipe.rs_push(r))
            pipe.launch_rs_flags(r)
            grads[r, w0:w1] = (xs[r].double().T @ dys[r].double()).float().reshape(-1).cpu()
        opt._next_slot = which
        opt._push_hyper()                   # the row holds the values for the step about to run (Adam: lr_t of t+1)
        pipe.launch_tail(which, which, opt.hyper_row(which).data_ptr())
        opt.iterations += 1
        pipe.launch_allgather()             # (the engine enqueues this at the start of the next step)
        for r in range(N):
            pipe.launch_param_wait(rank=r)  # every rank sees every peer's version flags
        torch.cuda.synchronize()
        ctx.check_error()
        for r in range(N):
            assert int(pipe._versions[r]) == step + 1
        mean = grads[0].clone()
        for r in range(1, N):
            mean += grads[r]
        cpu.grads.copy_(mean * (1.0 / N))
        cpu_opt.apply_gradients(cpu)
        reps = [buf.peer_tensor(r)[base:base + n].cpu() for r in range(N)]
        for rp in reps:
            assert torch.equal(rp, reps[0])                 # bit-identical replicas
        torch.testing.assert_close(reps[0][:w0], cpu.params[:w0], rtol=2e-5, atol=2e-6)
        torch.testing.assert_close(reps[0][w1:], cpu.params[w1:], rtol=2e-5, atol=2e-6)
        # the big gradient is a K=128 fp32-accumulated dot product: ~1e-6 relative error going into the update
        # (Adam divides by sqrt(v): elements whose gradient is ~0 amplify it, hence the looser absolute bound)
        torch.testing.assert_close(reps[0][w0:w1], cpu.params[w0:w1], rtol=1e-3, atol=2e-4)
        cpu.params.copy_(reps[0])                           # keep the oracle on the device's trajectory
    # optimizer state: small tensors replicated, the big tensor's rows sharded by owner -> gather and compare
    pipe.materialize_state()
    for k in range(opt.n_slots):
        got = arena.slots[k].cpu()
        torch.testing.assert_close(got[:w0], cpu.slots[k][:w0], rtol=2e-4, atol=1e-7)
        torch.testing.assert_close(got[w0:w1], cpu.slots[k][w0:w1], rtol=2e-2, atol=1e-5)


def test_param_wait_kernel_returns_when_target_reached():
    _need_gpu()
    from horovod_distributed_example_b200.parallel.arena import ParamArena
    from horovod_distributed_example_b200.parallel.peer import PeerContext
    from horovod_distributed_example_b200.parallel.pipeline import PipelinedStep
    ctx = PeerContext.emulate(2)
    arena = ParamArena(SPECS, "cuda", n_slots=2, double_grads=True, allocator=ctx.alloc_tensor)
    buf, base = ctx.buffer_of(arena.buffer)
    big = arena.spec("d1/k")
    pipe = PipelinedStep(ctx, buf, base, arena, 2, (0, big.offset), (big.offset, big.offset + big.numel),
                         (big.offset + big.numel, arena.n))
    pipe.launch_param_wait()           # version 0, flags 0: returns at once
    torch.cuda.synchronize()
    ctx.check_error()
