"""Write one SASS listing per hot kernel under profiles/sass/ (cuobjdump -sass of lib/libhvd_cuda.so).

    python tools/sass_listings.py

Each file is the kernel's complete SASS (addresses + mnemonics, encodings stripped) with a header that counts the
instructions proving the Blackwell-native / peer-memory paths: UTCHMMA (tcgen05.mma), LDTM (tcgen05.ld), UTMALDG /
UBLKCP (TMA tensor / bulk copies), LDGMC / STGMC-style multimem, *.SYS loads/stores/reductions (peer memory),
MEMBAR.SYS."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "horovod_distributed_example_b200", "lib", "libhvd_cuda.so")
OUT = os.path.join(ROOT, "profiles", "sass")

# (file name, substring of the demangled kernel name)
KERNELS = [
    ("tail_adadelta_nr8", "hvd_tail_kernel<2, 8>"),
    ("tail_adam_nr2", "hvd_tail_kernel<3, 2>"),
    ("allgather_tma", "hvd_allgather_tma_kernel"),
    ("allgather_ldst", "hvd_allgather_kernel("),
    ("rs_flags", "hvd_rs_flags_kernel"),
    ("ll_allreduce_nr8", "hvd_ll_allreduce_kernel<8>"),
    ("fused_nvls_adadelta_nr8", "hvd_fused_kernel<2, 2, 8>"),
    ("fused_twoshot_adadelta_nr8", "hvd_fused_kernel<1, 2, 8>"),
    ("fused_oneshot_adadelta_nr8", "hvd_fused_kernel<0, 2, 8>"),
    ("bcast", "hvd_bcast_kernel"),
    ("tc_gemm_bn128_mn_mn_dW_push", "tc_gemm_kernel<128, true, true>"),
    ("tc_gemm_bn128_k_mn_dense1_fwd", "tc_gemm_kernel<128, false, true>"),
    ("tc_gemm_bn64_k_k_dense1_dX_unpool", "tc_gemm_kernel<64, false, false>"),
    ("conv2_fprop", "conv2_fprop_kernel"),
    ("conv2_dgrad", "conv2_dgrad_kernel"),
    ("conv2_wgrad", "conv2_wgrad_kernel"),
    ("conv1_fwd", "conv1_fwd_kernel"),
    ("head", "head_kernel"),
]
COUNT = [("UTCHMMA", r"\bUTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"\bUTMALDG"), ("UBLKCP", r"\bUBLKCP"),
         ("UTCBAR", r"\bUTCBAR"), ("multimem (LDGMC/STG.MC/REDG.MC)", r"LDGMC|\.MC\b|MULTIMEM"),
         ("*.SYS memory ops", r"\b(LDG|STG|RED|REDG|ATOMG|LD|ST)\S*\.SYS"), ("MEMBAR.SYS", r"MEMBAR\S*\.SYS"),
         ("SYNCS (mbarrier)", r"\bSYNCS"), ("total instructions", r"^\s*/\*[0-9a-f]{4,}\*/")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"\n\s*Function : ", sass)
    funcs = {}
    for p in parts[1:]:
        mangled, _, body = p.partition("\n")
        dem = subprocess.run(["c++filt", mangled.strip()], capture_output=True, text=True).stdout.strip()
        funcs[dem] = body
    os.makedirs(OUT, exist_ok=True)
    for fname, key in KERNELS:
        hits = [d for d in funcs if key in d]
        if not hits:
            print("missing:", key, file=sys.stderr)
            continue
        dem = sorted(hits, key=len)[0]
        lines = []
        for ln in funcs[dem].splitlines():
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?)\s*/\*\s*0x[0-9a-f]+\s*\*/", ln)
            if m:
                lines.append(f"/*{m.group(1)}*/  {m.group(2).rstrip(' ;')};")
            elif ".headerflags" in ln or re.match(r"\s*\.L_", ln):
                lines.append(ln.rstrip())
        text = "\n".join(lines)
        hdr = [f"// {dem}", "// cuobjdump -sass lib/libhvd_cuda.so (sm_100a), encodings stripped", "//"]
        for label, pat in COUNT:
            n = len(re.findall(pat, text, flags=re.M))
            hdr.append(f"//   {label}: {n}")
        with open(os.path.join(OUT, fname + ".sass"), "w") as f:
            f.write("\n".join(hdr) + "\n\n" + text + "\n")
        print(fname, len(lines))


if __name__ == "__main__":
    main()
