"""Regenerate the tracked evidence under profiles/ from the built library and gpurun_out/ captures:
  * sass_summary.txt     - per-kernel counts of the SASS mnemonics that prove tcgen05 / TMA / NVLS use
  * ncu_<name>.csv       - selected `ncu --set full` metrics per captured launch
  * launches_step.csv    - every launch of one training step with its device time (ncu, cold cache)
"""
import csv
import glob
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
SO = os.path.join(ROOT, "horovod_distributed_example_b200", "lib", "libhvd_cuda.so")
MNEMONICS = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "LDGMC", "SYNCS",
             "RED.E.ADD.F32x4", "LDG.E.128.STRONG.SYS", "STG.E.128.STRONG.SYS", "HMMA"]
METRICS = ["Kernel Name", "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
           "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active"]


def sass_summary():
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    out, cur, counts = [], None, {}
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if cur:
                out.append((cur, counts))
            cur, counts = m.group(1), {}
            continue
        for mn in MNEMONICS:
            if mn in line and not (mn == "HMMA" and "UTCHMMA" in line):
                counts[mn] = counts.get(mn, 0) + 1
    if cur:
        out.append((cur, counts))
    demangled = subprocess.run(["cu++filt"] + [n for n, _ in out], capture_output=True, text=True).stdout.splitlines()
    with open(os.path.join(OUT, "sass_summary.txt"), "w") as f:
        f.write("# cuobjdump -sass lib/libhvd_cuda.so : mnemonic counts per kernel (sm_100a)\n")
        f.write("# UTCHMMA = tcgen05.mma, UTMALDG = TMA load, LDTM = tcgen05.ld, LDGMC = multimem.ld_reduce, "
                "*.STRONG.SYS = peer loads/stores\n")
        for (name, c), dn in zip(out, demangled):
            if c:
                f.write(f"{dn[:110]}\n    " + "  ".join(f"{k}={v}" for k, v in sorted(c.items())) + "\n")


def ncu_extract(rep, name):
    r = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True)
    rows = list(csv.reader(r.stdout.splitlines()))
    if len(rows) < 3:
        return
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(m) for m in METRICS if m in hdr]
    with open(os.path.join(OUT, f"ncu_{name}.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([hdr[i] for i in idx])
        w.writerow([units[i] for i in idx])
        for row in rows[2:]:
            w.writerow([row[i][:90] for i in idx])


def launches_from_report(rep):
    """launches_step.csv from a --set full report of one eager (HVD_NO_GRAPH=1) step."""
    r = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True)
    rows = list(csv.reader(r.stdout.splitlines()))
    if len(rows) < 3:
        return
    hdr, units = rows[0], rows[1]
    ki, ti = hdr.index("Kernel Name"), hdr.index("gpu__time_duration.sum")
    with open(os.path.join(OUT, "launches_step.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "device_time_us (ncu, cold cache, serialised)"])
        for row in rows[2:]:
            t = float(row[ti].replace(",", "")) / (1000.0 if units[ti] in ("ns", "nsecond") else 1.0)
            w.writerow([row[ki][:100], f"{t:.2f}"])


def launches():
    src = os.path.join(ROOT, "gpurun_out", "launches.csv")
    if not os.path.exists(src):
        reps = sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "prof_step*.ncu-rep")))
        if reps:
            launches_from_report(reps[-1])
        return
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    with open(os.path.join(OUT, "launches_step.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "device_time_us (ncu, cold cache, serialised)"])
        for r in rows[-14:]:
            t = float(r[vi].replace(",", "")) / (1000.0 if r[ui] == "ns" else 1.0)
            w.writerow([r[ki][:100], f"{t:.2f}"])


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    sass_summary()
    for rep in glob.glob(os.path.join(ROOT, "gpurun_out", "*.ncu-rep")):
        ncu_extract(rep, os.path.basename(rep).replace(".ncu-rep", ""))
    launches()
    print(sorted(os.listdir(OUT)))
