#!/usr/bin/env python
"""Coordinate descent over the static-schedule cost model (B200MD_COSTS) on the bench workload.  Run on a GPU box:
    python tools/tune_costs.py > gpurun_out/tune.log
Each evaluation is one short bench.py run (the schedule is built at topology creation, so one process per try)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ["pass", "chunk", "diaggroup", "diag", "bond", "angle", "dih", "gen", "p14"]


def run(costs):
    """one schedule (the force tasks carry no energy arithmetic any more: the fused integrator and the single-shot
    forces_kernel share it), timed on the bench workload"""
    env = dict(os.environ, B200MD_COSTS=",".join(str(int(c)) for c in costs))
    if "--forces" in sys.argv:  # the single-shot forces_kernel on BASELINE configs[1] instead of the fused integrator
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", "forces", "--no-cpu", "--no-energies", "--steps", "20", "--warmup", "3"]
    else:
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--no-cpu", "--no-extra", "--strong-total", "0", "--steps", "5",
               "--warmup", "3", "--inner", "500", "--energy-every", "0"]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True).stdout
    try:
        return json.loads(out.strip().splitlines()[-1])["value"]
    except Exception:
        return 0.0


def main():
    args = [x for x in sys.argv[1:] if not x.startswith("--")]
    default = "126,144,52,30,77,230,380,400,100"
    costs = [int(x) for x in (args[0] if args else default).split(",")]
    best = run(costs)
    print("start", costs, f"{best:.4e}", flush=True)
    for rnd in range(1 if "--forces" in sys.argv else 2):
        for k in range(len(costs)):
            if k == 7:
                continue  # no general torsions in ALA10
            for f in (0.7, 0.85, 1.2, 1.45):
                trial = list(costs)
                trial[k] = max(1, int(round(costs[k] * f)))
                v = run(trial)
                print(f"round {rnd} {NAMES[k]:9s} x{f:<4} {trial} {v:.4e}", flush=True)
                if v > best * 1.002:
                    best, costs = v, trial
        print("best so far", costs, f"{best:.4e}", flush=True)
    print("FINAL", ",".join(str(c) for c in costs), f"{best:.4e}")


if __name__ == "__main__":
    main()
