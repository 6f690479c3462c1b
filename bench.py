"""Headline benchmark: MNIST-CNN training images/sec (BASELINE.json metric), batch 128 per GPU.

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference|torch_nccl]

N>1 is launched by torchrun (one rank per GPU; RANK/LOCAL_RANK/WORLD_SIZE/MASTER_* from the
env).  Rank 0 prints ONE JSON line.  The model is the reference's CNN (mnist_keras.py:71-81)
with random-init weights, Adadelta with LR 1.0*size wrapped in DistributedOptimizer
(mnist_keras.py:84-87); data is synthetic 28x28x1 (no network for MNIST).

Timing: W warm-up steps, then K steps bracketed by barrier + cuda synchronize, timed with
CUDA events on the launching stream, max over ranks.  Inputs rotate through a device pool
larger than L2 (>= 160 MB) so no step re-reads its batch from cache.  The e2e number
goes through the public API (``model.train_on_batch_async``, each result awaited one step late) with
a pinned-host -> device copy of every batch and a device -> host read of the loss every step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 128


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch_nccl"])
    ap.add_argument("--optimizer", default="adadelta", choices=["adadelta", "adam"])
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region.

    The timed region of this benchmark is tens of milliseconds, shorter than nvidia-smi's fastest loop, so the
    primary sampler is an NVML polling thread (one sample every 5 ms, started right before the first timed step
    and stopped right after the last).  If NVML cannot be loaded it falls back to `nvidia-smi -lms 100`, whose
    samples then mostly fall just after the region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    PERIOD = 0.005

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []
        self.nvml = None
        self.samples = []           # (sm_mhz, reasons bitmask)
        self.max_mhz = None
        self._stop = threading.Event()

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            pr = torch.cuda.get_device_properties(self.idx)
            bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
        return pynvml, h

    def start(self):
        try:
            self.nvml, self.h = self._nvml_handle()
            self.max_mhz = float(self.nvml.nvmlDeviceGetMaxClockInfo(self.h, self.nvml.NVML_CLOCK_SM))
            self._sample()           # make sure the calls work before committing to this path
            self.thr = threading.Thread(target=self._poll, daemon=True)
            self.thr.start()
            return
        except Exception:
            self.nvml = None
            self.samples = []
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _sample(self):
        n = self.nvml
        self.samples.append((float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)),
                             int(n.nvmlDeviceGetCurrentClocksEventReasons(self.h))))

    def _poll(self):
        while not self._stop.wait(self.PERIOD):
            try:
                self._sample()
            except Exception:
                return

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.nvml is not None:
            self._stop.set()
            self.thr.join(1.0)
            n = self.nvml
            try:
                self._sample()       # one more right at the end of the region
            except Exception:
                pass
            bits = {"hw_slowdown": n.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": n.nvmlClocksEventReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": n.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": n.nvmlClocksEventReasonSwPowerCap,
                    "hw_power_brake_slowdown": n.nvmlClocksEventReasonHwPowerBrakeSlowdown}
            sm = sorted(s for s, _ in self.samples[1:]) or sorted(s for s, _ in self.samples)   # [0] predates the region
            seen = 0
            for _, r in self.samples[1:]:
                seen |= r
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(k for k, b in bits.items() if seen & b), "samples": len(sm), "source": "nvml, 5 ms period"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for nm, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi -lms 100"}


def reference_arm(args):
    """The reference cannot be installed here: /root/reference holds two scripts and a
    requirements.txt pinning tensorflow==1.15 + horovod==0.18.1 (no py3.12 wheels, no network,
    no MPI); there is no setup.py/pyproject to pip-install.  Recorded in DESIGN.md."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    why = "reference needs tensorflow==1.15 + horovod==0.18.1 + MPI (none installable offline on py3.12); " \
          "/root/reference has no installable package (pip install fails: no setup.py/pyproject.toml)"
    if os.path.isdir(ref_dir) and os.listdir(ref_dir):
        try:
            sys.path.insert(0, ref_dir)
            import horovod.keras  # noqa: F401
            import tensorflow  # noqa: F401
        except Exception as e:
            why = f"baseline/_ref present but not importable: {type(e).__name__}: {e}"
    print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


def make_model_ours(hvd, keras, opt_name):
    m = keras.Sequential([
        keras.layers.Conv2D(32, kernel_size=(3, 3), activation='relu', input_shape=(28, 28, 1)),
        keras.layers.Conv2D(64, (3, 3), activation='relu'),
        keras.layers.MaxPooling2D(pool_size=(2, 2)),
        keras.layers.Dropout(0.25),
        keras.layers.Flatten(),
        keras.layers.Dense(128, activation='relu'),
        keras.layers.Dropout(0.5),
        keras.layers.Dense(10, activation='softmax')], seed=1234)
    if opt_name == "adadelta":
        opt = keras.optimizers.Adadelta(1.0 * hvd.size())
    else:
        opt = keras.optimizers.Adam(0.001 * hvd.size())
    opt = hvd.DistributedOptimizer(opt)
    m.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=opt, metrics=['accuracy'])
    return m


def run_ours(args):
    import numpy as np
    import torch
    import horovod_distributed_example_b200 as hvd
    from horovod_distributed_example_b200 import keras

    hvd.init()
    n = hvd.size()
    if n != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but world size is {n}: launch with torchrun --nproc-per-node {args.gpus}")
    dev = hvd.device()
    if dev.type != "cuda":
        raise SystemExit("bench.py needs a CUDA device (the CPU oracle is not a benchmark target)")
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    model = make_model_ours(hvd, keras, args.optimizer)
    rng = np.random.default_rng(1000 + hvd.rank())
    pool_n = 416                                  # 416 * 128 * 784 * 4 B = 167 MB > 126 MB L2
    xs_host = torch.from_numpy(rng.random((pool_n, BATCH, 28, 28, 1), dtype=np.float32)).pin_memory()
    ys_host = torch.from_numpy(rng.integers(0, 10, (pool_n, BATCH)).astype(np.int64)).pin_memory()
    xs_dev, ys_dev = xs_host.to(dev), ys_host.to(dev)
    model._ensure_built((BATCH, 28, 28, 1))
    hvd.broadcast_global_variables(0, model)
    be = model._backend
    launches_per_step = getattr(be, "launches_per_step", None)
    if launches_per_step is None:          # torch-autograd backend: only the fused update kernels are ours
        fused = getattr(model.optimizer, "_fused", None)
        launches_per_step = len(fused.buckets) if fused is not None else 0

    xh_np, yh_np = xs_host.numpy(), ys_host.numpy()
    host_fed = hasattr(be, "train_host_lazy")

    def step(i):
        if host_fed:
            # the engine's own staging: batch i is copied pinned-host -> device on the copy stream into the
            # input buffers of the next step parity while step i-1 computes; returns device scalars (no sync)
            return be.train_host_lazy(xh_np[i % pool_n], yh_np[i % pool_n])
        return be.train_step(xs_dev[i % pool_n], ys_dev[i % pool_n])

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    hvd.barrier()
    sampler = ClockSampler(dev.index)
    if hvd.rank() == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(args.steps):
        loss, _ = step(args.warmup + i)
    e1.record()
    torch.cuda.synchronize()
    hvd.barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if hvd.rank() == 0 else None
    ms_t = torch.tensor([ms], dtype=torch.float64)
    all_ms = hvd.allgather(ms_t)
    ms_max = float(all_ms.max())
    final_loss = float(loss)

    e2e = None
    if not args.no_e2e:
        xh = xs_host.numpy()
        yh = ys_host.numpy()
        for i in range(max(3, args.warmup // 2)):
            model.train_on_batch(xh[i % pool_n], yh[i % pool_n])
        torch.cuda.synchronize()
        hvd.barrier()
        # Public API, pipelined by one step: every step copies ITS batch from pinned host memory to the
        # device and ITS (loss, accuracy, step) lands in pinned host memory (written by a kernel of the
        # step's backward, 12 bytes); the result of step t is awaited while step t+1 is already enqueued, so the H2D
        # copy of batch t+1 overlaps the compute of step t.
        t0 = time.perf_counter()
        pending = None
        for i in range(args.steps):
            h = model.train_on_batch_async(xh[(7 + i) % pool_n], yh[(7 + i) % pool_n])
            if pending is not None:
                out = pending.result()
            pending = h
        out = pending.result()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        hvd.barrier()
        e2e_ms = float(hvd.allgather(torch.tensor([(t1 - t0) * 1e3], dtype=torch.float64)).max())
        e2e = {"value": BATCH * n * args.steps / (e2e_ms / 1e3), "unit": "images/sec",
               "h2d_bytes_per_step": BATCH * 28 * 28 * 4 + BATCH * 8,
               "d2h_bytes_per_step": 12, "ms_per_step": e2e_ms / args.steps, "last_loss": float(out[0]),
               "api": "Model.train_on_batch_async (result awaited one step late)"}

    if hvd.rank() == 0:
        val = BATCH * n * args.steps / (ms_max / 1e3)
        gl = (launches_per_step or 0) * args.steps
        line = {"metric": "MNIST-CNN training images/sec", "value": val, "unit": "images/sec",
                "n_gpus": n, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "tf32 (fp32 storage, fp32 accumulate/optimizer)",
                "data": "synthetic 28x28x1 (MNIST shape), random-init weights", "impl": "ours",
                "config": {"model": "mnist_keras.py CNN (conv3x3 32/64, maxpool, dropout, dense 128/10)",
                           "global_batch": BATCH * n, "batch_per_gpu": BATCH, "seq_len": None,
                           "parallelism": f"dp{n}", "optimizer": f"{args.optimizer} lr={'1.0' if args.optimizer == 'adadelta' else '0.001'}*size",
                           "backend": getattr(be, "name", type(be).__name__),
                           "allreduce": getattr(getattr(model.optimizer, "_fused", None), "describe", lambda: ["torch/nccl"])(),
                           "l2": ("inputs rotate through a 167 MB pinned host pool (> 126 MB L2), H2D on a copy stream under the previous step"
                                  if host_fed else "inputs rotate through a 167 MB device pool (> 126 MB L2)")},
                "clocks": clocks, "e2e": e2e, "gpu_launches": gl, "final_loss": final_loss}
        print(json.dumps(line))
    hvd.shutdown()
    return 0


def run_torch_nccl(args):
    """Stock baseline of record: the same CNN in eager PyTorch
    (cuDNN/cuBLAS, TF32 allowed), gradients averaged with torch.distributed NCCL all_reduce on one
    flattened bucket, torch.optim Adadelta/Adam.  None of this repo's kernels or engine."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import torch.nn as nn
    import torch.nn.functional as F

    rank = int(os.environ.get("RANK", 0))
    n = int(os.environ.get("WORLD_SIZE", 1))
    lr_ = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr_)
    dev = torch.device("cuda", lr_)
    if n > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        dist.init_process_group("nccl", rank=rank, world_size=n, device_id=dev)
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.benchmark = True

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.c1, self.c2 = nn.Conv2d(1, 32, 3), nn.Conv2d(32, 64, 3)
            self.f1, self.f2 = nn.Linear(9216, 128), nn.Linear(128, 10)

        def forward(self, x):
            x = F.relu(self.c1(x))
            x = F.max_pool2d(F.relu(self.c2(x)), 2)
            x = F.dropout(x, 0.25, self.training).flatten(1)
            x = F.dropout(F.relu(self.f1(x)), 0.5, self.training)
            return self.f2(x)

    torch.manual_seed(1234)
    net = Net().to(dev).to(memory_format=torch.channels_last)
    params = list(net.parameters())
    if args.optimizer == "adadelta":
        opt = torch.optim.Adadelta(params, lr=1.0 * n, rho=0.95, eps=1e-7)
    else:
        opt = torch.optim.Adam(params, lr=0.001 * n, eps=1e-7)
    rng = np.random.default_rng(1000 + rank)
    pool_n = 416
    xs_host = torch.from_numpy(rng.random((pool_n, BATCH, 1, 28, 28), dtype=np.float32)).pin_memory()
    ys_host = torch.from_numpy(rng.integers(0, 10, (pool_n, BATCH)).astype(np.int64)).pin_memory()
    xs_dev, ys_dev = xs_host.to(dev), ys_host.to(dev)

    def step(x, y):
        opt.zero_grad(set_to_none=False)
        loss = F.cross_entropy(net(x), y)
        loss.backward()
        if n > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            flat.div_(n)
            off = 0
            for p in params:
                k = p.numel()
                p.grad.copy_(flat[off:off + k].view_as(p))
                off += k
        opt.step()
        return loss

    net.train()
    for i in range(args.warmup):
        step(xs_dev[i % pool_n], ys_dev[i % pool_n])
    torch.cuda.synchronize()
    if n > 1:
        dist.barrier()
    sampler = ClockSampler(lr_)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss = step(xs_dev[(args.warmup + i) % pool_n], ys_dev[(args.warmup + i) % pool_n])
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if n > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    clocks = sampler.stop() if rank == 0 else None
    # e2e: pinned host batch -> device every step, loss.item() every step
    t0 = time.perf_counter()
    for i in range(args.steps):
        x = xs_host[(7 + i) % pool_n].to(dev, non_blocking=True)
        y = ys_host[(7 + i) % pool_n].to(dev, non_blocking=True)
        step(x, y).item()
    torch.cuda.synchronize()
    e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3], device=dev, dtype=torch.float64)
    if n > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms_max = float(ms)
        print(json.dumps({
            "metric": "MNIST-CNN training images/sec", "value": BATCH * n * args.steps / (ms_max / 1e3),
            "unit": "images/sec", "n_gpus": n, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "tf32 (cuDNN/cuBLAS allow_tf32)", "data": "synthetic 28x28x1",
            "impl": "torch_nccl",
            "config": {"model": "mnist_keras.py CNN, eager PyTorch + NCCL all_reduce + torch.optim",
                       "global_batch": BATCH * n, "parallelism": f"dp{n}", "optimizer": args.optimizer},
            "clocks": clocks, "gpu_launches": 0,
            "e2e": {"value": BATCH * n * args.steps / (float(e2e_ms) / 1e3), "unit": "images/sec",
                    "h2d_bytes_per_step": BATCH * 784 * 4 + BATCH * 8, "d2h_bytes_per_step": 4}}))
    if n > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)
    if args.impl == "torch_nccl":
        return run_torch_nccl(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
