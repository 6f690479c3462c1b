"""Run `.ps_project/workflow.yaml` locally: launch the experiment through hvdrun and evaluate the
acceptance checks on the scalars the run recorded.  This is the in-repo equivalent of the reference's
Gradient multi-node experiment plus its GradientCI metric check (loss must average 0.0..0.3).

    python tools/run_workflow.py [--workflow .ps_project/workflow.yaml] [--np N] [--env K=V ...]

Exit code 0 iff the experiment exits 0 and every check passes.
"""
import argparse
import json
import os
import subprocess
import sys

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# scalar tags a check on "<name>" may be satisfied by, most specific first
TAGS = {"loss": ("epoch_loss", "batch_loss", "loss"), "accuracy": ("epoch_accuracy", "batch_accuracy", "accuracy")}
AGG = {"mean": lambda v: sum(v) / len(v), "min": min, "max": max, "last": lambda v: v[-1]}


def read_scalars(path):
    out = {}
    if os.path.exists(path):
        with open(path) as f:
            for line in f:
                rec = json.loads(line)
                out.setdefault(rec["tag"], []).append(rec["value"])
    return out


def run_checks(scalars, checks):
    results = []
    for chk in checks or []:
        series = next((scalars[t] for t in TAGS.get(chk["metric"], (chk["metric"],)) if scalars.get(t)), None)
        if not series:
            results.append({"metric": chk["metric"], "passed": False, "reason": "no values recorded"})
            continue
        value = AGG[chk.get("aggregate", "mean")](series)
        results.append({"metric": chk["metric"], "aggregate": chk.get("aggregate", "mean"), "value": value,
                        "range": [chk["min"], chk["max"]], "passed": chk["min"] <= value <= chk["max"]})
    return results


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--workflow", default=os.path.join(ROOT, ".ps_project", "workflow.yaml"))
    ap.add_argument("--np", type=int, default=None, help="override the experiment's process count")
    ap.add_argument("--env", action="append", default=[])
    a = ap.parse_args(argv)
    wf = yaml.safe_load(open(a.workflow))
    exp = yaml.safe_load(open(os.path.join(ROOT, wf["run"]["experiment"])))
    metrics_path = os.path.join(ROOT, "metrics.jsonl")
    if os.path.exists(metrics_path):
        os.remove(metrics_path)
    cmd = [sys.executable, "-m", "horovod_distributed_example_b200.run", "-np", str(a.np or exp["processes"])]
    for item in exp.get("export", []):
        cmd += ["-x", item]
    cmd.append(exp["script"])
    env = dict(os.environ, HVD_METRICS_PATH=metrics_path)
    env.update(dict(e.split("=", 1) for e in a.env))
    print(f"[{wf['name']}] {' '.join(cmd)}", flush=True)
    rc = subprocess.run(cmd, cwd=ROOT, env=env).returncode
    ok = rc == 0
    if rc != 0:
        print(f"[{wf['name']}] experiment failed with exit code {rc}")
    for res in run_checks(read_scalars(metrics_path), wf.get("accept")):
        print(f"[{wf['name']}] check {json.dumps(res)}")
        ok = ok and res["passed"]
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
