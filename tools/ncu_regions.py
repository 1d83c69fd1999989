#!/usr/bin/env python
"""Region-level attribution of an ncu source page: groups tools/ncu_by_line.py rows by the device function
(line ranges of b200md_forces.cuh parsed from the file itself) the SASS instruction was inlined from."""
import collections
import csv
import io
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ncu_by_line as n  # noqa: E402

ROOT = n.ROOT
SRC = os.path.join(ROOT, "molecular_dynamics_b200", "csrc", "b200md_forces.cuh")


def function_ranges():
    names, lines = [], open(SRC).read().splitlines()
    for i, l in enumerate(lines, 1):
        m = re.match(r"__device__ __forceinline__ \w+ (\w+)\(", l)
        if m:
            names.append((i, m.group(1)))
    out = []
    for k, (start, name) in enumerate(names):
        end = names[k + 1][0] - 3 if k + 1 < len(names) else len(lines)
        out.append((name, start - 2, end))
    return out


def main():
    report, kernel = sys.argv[1], sys.argv[2]
    sass = n.sass_lines(kernel)
    raw = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    col = {h: i for i, h in enumerate(hdr)}
    regions = function_ranges()
    agg = collections.defaultdict(collections.Counter)
    tot = collections.Counter()
    last = "other"
    for (loc, ins), r in zip(sass, body):
        f, l = loc
        if f == "b200md_forces.cuh":
            reg = next((nm for nm, a, b in regions if a <= l <= b), "forces_other")
            last = reg
        elif f == "b200md.cu":
            reg = last = "kernel_body"
        else:
            reg = last + " (inlined helpers)"
        op = ins.split()[0] if not ins.startswith("@") else ins.split()[1]
        if op.startswith(("LDL", "STL")):
            reg = "local-memory spills"
        for k, name in (("inst", "Instructions Executed"), ("smp", "# Samples"), ("wave", "L1 Wavefronts Shared")):
            v = float(r[col[name]] or 0)
            agg[reg][k] += v
            tot[k] += v
    print(f"warp instructions {tot['inst']:.4e}  samples {tot['smp']:.0f}  smem wavefronts {tot['wave']:.4e}")
    print(f"{'region':40s} {'inst%':>7s} {'samples%':>9s} {'smem wave%':>10s}")
    for reg, d in sorted(agg.items(), key=lambda kv: -kv[1]["smp"]):
        print(f"{reg:40s} {100 * d['inst'] / tot['inst']:7.2f} {100 * d['smp'] / tot['smp']:9.2f} {100 * d['wave'] / max(tot['wave'], 1):10.2f}")


if __name__ == "__main__":
    main()
