"""Stand-alone device time of the conv2 backward kernels at different CTA counts (warm, back-to-back launches).
Shows whether a kernel is bound by its tensor-core issue time (scales with tiles per CTA) or by something else."""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from horovod_distributed_example_b200.ops import native

B = 128


def timed(fn, n=40):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) * 1e3 / n, 2)


def main():
    lib = native.cuda_lib(required=True)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.hvd_conv2_dgrad.argtypes = [vp] * 4 + [ci, ci] + [vp] * 4
    lib.hvd_conv2_wgrad.argtypes = [vp] * 4 + [ci, ci, vp]
    x = torch.rand(B, 28, 28, device="cuda")
    a1 = torch.relu(torch.randn(B, 26, 26, 32, device="cuda"))
    w2 = torch.randn(3, 3, 32, 64, device="cuda") * 0.05
    dy2 = torch.randn(B, 24, 24, 64, device="cuda")
    gw1, gb1 = torch.zeros(9, 32, device="cuda"), torch.zeros(32, device="cuda")
    gw, gb = torch.zeros(3, 3, 32, 64, device="cuda"), torch.zeros(64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    out = {"dgrad_fused_w1": {}, "dgrad_plain": {}, "wgrad": {}}
    da1 = torch.empty_like(a1)
    for ctas in (148, 128, 110, 96, 64):
        out["dgrad_fused_w1"][ctas] = timed(lambda: lib.hvd_conv2_dgrad(dy2.data_ptr(), w2.data_ptr(), a1.data_ptr(), None, B, ctas, s,
                                                                       x.data_ptr(), gw1.data_ptr(), gb1.data_ptr()))
        out["dgrad_plain"][ctas] = timed(lambda: lib.hvd_conv2_dgrad(dy2.data_ptr(), w2.data_ptr(), a1.data_ptr(), da1.data_ptr(), B, ctas, s,
                                                                    None, None, None))
    for ctas in (128, 96, 64, 48, 38):
        out["wgrad"][ctas] = timed(lambda: lib.hvd_conv2_wgrad(a1.data_ptr(), dy2.data_ptr(), gw.data_ptr(), gb.data_ptr(), B, ctas, s))
    # in-kernel phase timers of CTA 0 (tile index 2 of its range), dgrad with the fused conv1 wgrad at 110 CTAs
    lib.hvd_conv2_dgrad(dy2.data_ptr(), w2.data_ptr(), a1.data_ptr(), None, B, 110, s, x.data_ptr(), gw1.data_ptr(), gb1.data_ptr())
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 64)()
    lib.hvd_conv2_bwd_debug_read.argtypes = [ctypes.c_void_p]
    lib.hvd_conv2_bwd_debug_read(buf)
    t = list(buf)
    base = t[0]
    out["dgrad_epilogue_cycles(tile2)"] = {"wait_acc": t[1] - t[0], "tmem_ld": t[6] - t[1], "xch_store+shuffles": t[7] - t[6], "bar+fixup": t[2] - t[7], "a1_mask": t[3] - t[2],
                                            "wait_stage_free": t[4] - t[3], "staging": t[5] - t[4], "total": t[5] - t[0]}
    for k, o in (("mma_tile2", 8), ("mma_tile3", 16)):
        out[f"dgrad_{k}_cycles"] = {"start_rel": t[o] - base, "wait_acc_empty": t[o + 1] - t[o], "half0_ready_rel": t[o + 4] - t[o + 1],
                                    "half1_ready_rel": t[o + 5] - t[o + 1], "main_issued": t[o + 2] - t[o + 1], "w1_issue": t[o + 3] - t[o + 2]}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
