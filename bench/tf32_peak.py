"""Measure the box's TF32 GEMM throughput (cuBLAS through torch.matmul, fp32 storage, allow_tf32) - the roofline
denominator for the tcgen05 kind::tf32 kernels (profiles/README.md).  MEASURED_PEAKS.json only carries bf16.

    python bench/tf32_peak.py  ->  one JSON line: burst (best of 10) and sustained (back to back for ~3 s)"""
import json
import time

import torch


def main():
    torch.backends.cuda.matmul.allow_tf32 = True
    n = 8192
    a = torch.randn(n, n, device="cuda")
    b = torch.randn(n, n, device="cuda")
    c = torch.empty(n, n, device="cuda")
    for _ in range(3):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    flop = 2.0 * n ** 3
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    iters = 0
    e0.record()
    while time.time() - t0 < 3.0:
        for _ in range(20):
            torch.matmul(a, b, out=c)
        iters += 20
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    sustained_ms = e0.elapsed_time(e1) / iters
    # bf16 on the same box, same method, for the ratio
    ab, bb = a.bfloat16(), b.bfloat16()
    cb = torch.empty(n, n, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        torch.matmul(ab, bb, out=cb)
    torch.cuda.synchronize()
    bestb = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(ab, bb, out=cb)
        e1.record()
        torch.cuda.synchronize()
        bestb = min(bestb, e0.elapsed_time(e1))
    print(json.dumps({"tf32_tflops_burst": flop / best / 1e9, "tf32_tflops_sustained": flop / sustained_ms / 1e9,
                      "bf16_tflops_burst": flop / bestb / 1e9, "n": n, "gpu": torch.cuda.get_device_name(0),
                      "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS), CUDA events; burst = best of 10, "
                             "sustained = back to back for 3 s"}))


if __name__ == "__main__":
    main()
