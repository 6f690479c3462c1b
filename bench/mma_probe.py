"""Cycles per tcgen05.mma (cta_group::1, kind::tf32, 128 x N x 8) as a function of N and of how many
independent accumulators the K loop rotates over.  Writes gpurun_out/mma_probe.json."""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from horovod_distributed_example_b200.ops import native


def main():
    lib = native.cuda_lib(required=True)
    lib.hvd_mma_probe.restype = ctypes.c_int
    lib.hvd_mma_probe.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_void_p]
    rows = []
    iters = 1152
    for ctas in (1, 148):
        out = torch.zeros(ctas, dtype=torch.int64, device="cuda")
        for M in (128, 64):
            for N in (16, 32, 64, 96, 128, 192, 256):
                if M == 64 and N == 16:
                    continue
                for n_acc in (1, 2):
                    if n_acc * N > 512:
                        continue
                    for _ in range(2):
                        rc = lib.hvd_mma_probe((M << 16) | N, iters, n_acc, 0, ctas, out.data_ptr(), None)
                        assert rc == 0, rc
                        torch.cuda.synchronize()
                    cyc = float(out.max()) / iters
                    flop = 2 * M * N * 8
                    rows.append({"ctas": ctas, "M": M, "N": N, "n_acc": n_acc, "cycles_per_mma": round(cyc, 1),
                                 "flop_per_clk_per_sm": round(flop / cyc, 0)})
                    print(rows[-1], flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/mma_probe.json", "w"), indent=1)


if __name__ == "__main__":
    main()
