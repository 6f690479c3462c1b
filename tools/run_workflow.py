"""Run a `.ps_project/config.yaml` workflow locally: launch each step's master command and evaluate
the metric checks on what the run recorded - the stand-in for Gradient's `experiment.run_mpi_multi_node`
+ GradientCI check (reference .ps_project/config.yaml:5-11, distributed-keras-sample.yaml:8).

    python tools/run_workflow.py [--config .ps_project/config.yaml] [--np N] [--env K=V ...]

Exit code 0 iff every step exits 0 and every check passes.
"""
import argparse
import json
import os
import shlex
import subprocess
import sys

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def evaluate(metrics_path, checks):
    vals = {}
    if os.path.exists(metrics_path):
        for line in open(metrics_path):
            r = json.loads(line)
            vals.setdefault(r["tag"], []).append(r["value"])
    alias = {"tensorflow:loss": ("epoch_loss", "batch_loss", "loss")}
    out = []
    for metric, spec in (checks or {}).items():
        lo, hi = (float(x) for x in spec["target"].split(".."))
        series = next((vals[t] for t in alias.get(metric, (metric,)) if vals.get(t)), [])
        if not series:
            out.append({"metric": metric, "passed": False, "reason": "no values recorded"})
            continue
        agg = {"mean": sum(series) / len(series), "min": min(series), "max": max(series),
               "last": series[-1]}[spec.get("aggregate", "mean")]
        out.append({"metric": metric, "aggregate": spec.get("aggregate", "mean"), "value": agg,
                    "target": spec["target"], "passed": lo <= agg <= hi})
    return out


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default=os.path.join(ROOT, ".ps_project", "config.yaml"))
    ap.add_argument("--np", type=int, default=None, help="override the process count in the master command")
    ap.add_argument("--env", action="append", default=[])
    a = ap.parse_args(argv)
    cfg = yaml.safe_load(open(a.config))
    ok = True
    for wname, wf in cfg.get("workflows", {}).items():
        metrics_path = os.path.join(ROOT, "metrics.jsonl")
        if os.path.exists(metrics_path):
            os.remove(metrics_path)
        for step in wf.get("steps", []):
            params = yaml.safe_load(open(os.path.join(ROOT, step["paramsFile"])))
            cmd = shlex.split(params["masterCommand"])
            if a.np is not None and "-np" in cmd:
                cmd[cmd.index("-np") + 1] = str(a.np)
            env = dict(os.environ, HVD_METRICS_PATH=metrics_path)
            env.update(dict(e.split("=", 1) for e in a.env))
            print(f"[workflow {wname}] step {step['name']}: {' '.join(cmd)}", flush=True)
            rc = subprocess.run(cmd, cwd=ROOT, env=env).returncode
            if rc != 0:
                print(f"[workflow {wname}] step failed with exit code {rc}")
                ok = False
        for res in evaluate(metrics_path, wf.get("checks")):
            print(f"[workflow {wname}] check {json.dumps(res)}")
            ok = ok and res["passed"]
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
