"""Gradient-allreduce latency / bus-bandwidth sweep (BASELINE.json config 4): the peer-memory kernels
(one-shot, two-shot, NVLS) against ``ncclAllReduce`` through torch.distributed, 1 KB - 1 GB.

Launch:  torchrun --nproc-per-node N bench/allreduce_sweep.py [--max-mb 1024]
Every number is timed with CUDA events on the launching stream and is the MAX over ranks.
busbw = (S / t) * 2 (N-1) / N   (NCCL convention).
Writes gpurun_out/allreduce_sweep_N<N>.json and the crossover table profiles/variant_table.json reads.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200.parallel import peer as P
from horovod_distributed_example_b200.parallel import variants as V


def time_fn(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    hvd.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], dtype=torch.float64)
    return float(hvd.allgather(t).max())          # us, max over ranks


def time_graph(fn, reps=20, iters=10):
    """Device-side latency: `reps` back-to-back launches captured in one CUDA graph, so the host
    launch rate (about 10 us per ctypes/NCCL call) does not floor the small-message numbers."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    hvd.barrier()
    g = torch.cuda.CUDAGraph()
    cap = torch.cuda.Stream()
    cap.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(g, stream=cap):
        for _ in range(reps):
            fn()
    torch.cuda.current_stream().wait_stream(cap)
    g.replay()
    torch.cuda.synchronize()
    hvd.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (iters * reps)], dtype=torch.float64)
    return float(hvd.allgather(t).max())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    a = ap.parse_args()
    hvd.init()
    n, r = hvd.size(), hvd.rank()
    ctx = P.context()
    max_bytes = a.max_mb << 20
    buf = ctx.alloc(2 * max_bytes)
    buf.tensor.normal_()
    nccl_t = torch.randn(max_bytes // 4, device="cuda")
    has_nvls = bool(buf.mc_ptr)
    rows = []
    size = 1024
    while size <= max_bytes:
        nel = size // 4
        iters = 200 if size <= (1 << 20) else (50 if size <= (64 << 20) else 10)
        row = {"bytes": size, "timing": "cuda-graph of 20 launches" if size <= (4 << 20) else "event loop"}

        def kern(variant, ctas):
            args = P.HvdFusedArgs()
            args.start, args.end = 0, nel
            args.grad_off = 0
            args.out_off = (max_bytes // 4) if variant == P.ONESHOT else 0     # one-shot must not run in place
            args.scale, args.kind, args.variant, args.slot, args.exit_barrier = 1.0 / n, 0, variant, 7, 1
            return lambda: ctx.launch_fused(buf, args, ctas)

        for name, variant in (("oneshot", P.ONESHOT), ("twoshot", P.TWOSHOT), ("nvls", P.NVLS)):
            if variant == P.NVLS and not has_nvls:
                continue
            if variant == P.ONESHOT and size > (32 << 20):
                continue
            best = None
            for ctas in sorted({V.default_ctas(size, variant, n), 8, 32, 148}):
                if ctas > 148:
                    continue
                t = time_graph(kern(variant, ctas)) if size <= (4 << 20) else time_fn(kern(variant, ctas), iters)
                if best is None or t < best[0]:
                    best = (t, ctas)
            row[name + "_us"], row[name + "_ctas"] = best
            row[name + "_busbw_GBs"] = size / best[0] * 1e-3 * 2 * (n - 1) / n
        if size <= P.LL_MAX_BYTES:
            ll_fn = lambda: ctx.launch_ll(buf, 0, max_bytes // 4, nel, 1.0 / n)
            row["ll_us"] = time_graph(ll_fn)
            row["ll_ctas"] = (nel // 2 + 255) // 256
            row["ll_busbw_GBs"] = size / row["ll_us"] * 1e-3 * 2 * (n - 1) / n
        nccl_fn = lambda: dist.all_reduce(nccl_t[:nel])
        t = time_graph(nccl_fn) if size <= (4 << 20) else time_fn(nccl_fn, iters)
        row["nccl_us"] = t
        row["nccl_busbw_GBs"] = size / t * 1e-3 * 2 * (n - 1) / n
        rows.append(row)
        if r == 0:
            print(json.dumps(row), flush=True)
        size *= 4
    ctx.check_error()
    if r == 0:
        # crossover table: largest size where one-shot is still the fastest of ours; smallest where NVLS wins
        one_max, nvls_min, cta_tab = 0, None, []
        for row in rows:
            ours = {k: row[k + "_us"] for k in ("oneshot", "twoshot", "nvls") if k + "_us" in row}   # (LL is picked by size)
            best = min(ours, key=ours.get)
            if best == "oneshot":
                one_max = row["bytes"]
            if best == "nvls" and nvls_min is None:
                nvls_min = row["bytes"]
            cta_tab.append([row["bytes"], row[best + "_ctas"]])
        out = {"n_gpus": n, "nvls": has_nvls, "rows": rows,
               "table": {str(n): {"oneshot_max_bytes": one_max, "nvls_min_bytes": nvls_min or (1 << 62),      # never won: never picked
                                  "ctas": cta_tab}}}
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"allreduce_sweep_N{n}.json"), "w"), indent=1)
    hvd.shutdown()


if __name__ == "__main__":
    main()
