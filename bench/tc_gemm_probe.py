"""Bring-up probe for the tcgen05 GEMM: runs each major-ness combination in its own process
(a bad descriptor traps and poisons the CUDA context) and, when one mismatches, sweeps LBO/SBO
candidates so a single GPU call tells us which encoding the hardware wants."""
import itertools
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(a_mn, b_mn, bn, M, N, K, ks, dbg):
    import torch
    from horovod_distributed_example_b200.ops.gemm import tc_gemm, EPI_ATOMIC_ADD, EPI_STORE
    torch.manual_seed(0)
    A = torch.randn(M, K, device="cuda")
    B = torch.randn(K, N, device="cuda")
    ref = (A.double() @ B.double()).float()
    Ain = A.t().contiguous() if a_mn else A
    Bin = B if b_mn else B.t().contiguous()
    out = tc_gemm(Ain, Bin, M, N, K, a_mn, b_mn, bn, ks, EPI_ATOMIC_ADD if ks > 1 else EPI_STORE, dbg=dbg)
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    if os.environ.get("PROBE_VERBOSE"):
        print("out absmax", out.abs().max().item(), "zeros", (out == 0).float().mean().item(), "nan", out.isnan().any().item())
    rel = err / ref.abs().max().item()
    print(json.dumps({"a_mn": a_mn, "b_mn": b_mn, "bn": bn, "M": M, "N": N, "K": K, "ks": ks, "dbg": dbg,
                      "max_abs_err": err, "rel": rel, "ok": rel < 5e-3}))


def sub(args):
    r = subprocess.run([sys.executable, __file__, "one"] + [str(a) for a in args], capture_output=True, text=True,
                       timeout=120)
    line = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if line:
        return json.loads(line[-1])
    return {"args": args, "crash": (r.stderr or r.stdout)[-300:]}


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "one":
        v = [int(x) for x in sys.argv[2:]]
        one(v[0], v[1], v[2], v[3], v[4], v[5], v[6], tuple(v[7:11]))
        sys.exit(0)
    results = []
    for a_mn, b_mn in itertools.product((0, 1), (0, 1)):
        res = sub([a_mn, b_mn, 128, 256, 256, 128, 1, 0, 0, 0, 0])
        results.append(res)
        print(res, flush=True)
        if not res.get("ok"):
            for lbo, sbo in [(4096, 1024), (512, 4096), (4096, 128), (128, 4096), (1024, 512), (16, 512)]:
                dbg = [lbo if a_mn else 0, sbo if a_mn else 0, lbo if b_mn else 0, sbo if b_mn else 0]
                r2 = sub([a_mn, b_mn, 128, 256, 256, 128, 1] + dbg)
                results.append(r2)
                print("  sweep", r2, flush=True)
    # shapes used by the model: dense1 fwd/dX/dW, with N tile 64 and split-K
    for args in ([0, 1, 128, 128, 128, 9216, 72, 0, 0, 0, 0], [0, 0, 128, 128, 9216, 128, 1, 0, 0, 0, 0],
                 [1, 1, 128, 9216, 128, 128, 1, 0, 0, 0, 0], [0, 1, 64, 384, 64, 288, 1, 0, 0, 0, 0],
                 [0, 0, 64, 200, 100, 64, 1, 0, 0, 0, 0]):
        res = sub(args)
        results.append(res)
        print(res, flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(results, open(os.path.join(ROOT, "gpurun_out", "tc_gemm_probe.json"), "w"), indent=1)
