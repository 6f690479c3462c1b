"""Runs csrc/kernels/desc_probe.cu over shifted starts / strides / base_offset values."""
import ctypes
import itertools
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from horovod_distributed_example_b200.ops import native

lib = native.cuda_lib()
vp, ci, cu = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint
lib.hvd_desc_probe.restype = ci
lib.hvd_desc_probe.argtypes = [vp, vp, vp, ci, ci, ci, cu, cu, ci, vp]
torch.manual_seed(0)
rows = 224
Ag = torch.randn(rows, 32, device="cuda")
Bg = torch.randn(64, 32, device="cuda")
out = []


def run(a_mn, shift, lbo, sbo, bo):
    D = torch.zeros(128, 64, device="cuda")
    rc = lib.hvd_desc_probe(Ag.data_ptr(), Bg.data_ptr(), D.data_ptr(), a_mn, rows, shift, lbo, sbo, bo,
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    m = torch.arange(128, device="cuda")
    if not a_mn:
        ridx = shift + (m // 8) * (sbo // 128) + m % 8
        A = Ag[ridx]                                     # [128][32]
    else:
        k = torch.arange(32, device="cuda")
        ridx = shift + k[None, :] + ((m // 32) * (lbo // 128))[:, None]      # [128][32] row per (m,k)
        A = Ag[ridx, (m % 32)[:, None].expand(128, 32)]
    ref = A.double() @ Bg.double().t()
    rel = ((D.double() - ref).abs().max() / ref.abs().max()).item()
    res = {"a_mn": a_mn, "shift": shift, "lbo": lbo, "sbo": sbo, "base_offset": bo, "rc": rc, "rel": rel, "ok": rel < 5e-3}
    out.append(res)
    print(res, flush=True)


for shift, sbo in itertools.product((0, 1, 2, 3, 5, 8, 10), (1024, 1280)):
    for bo in sorted({0, shift & 7}):
        run(0, shift, 16, sbo, bo)
for shift, lbo in itertools.product((0, 1, 2, 3, 4, 10, 11), (128, 4096)):
    for bo in sorted({0, shift & 3, shift & 7}):
        run(1, shift, lbo, 512, bo)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "desc_probe.json"), "w"), indent=1)
