#!/usr/bin/env python
"""Attribute an `ncu --page source --csv` dump to CUDA source lines.

ncu's CSV page lists SASS instructions in program order; nvdisasm --print-line-info lists the same
instructions with `//## File "...", line N` markers.  Zipping the two gives per-source-line totals of
executed instructions, stall samples and shared-memory wavefronts (read here, where there is no GPU).

usage: tools/ncu_by_line.py prof.ncu-rep <kernel-substring> [--top 40]
"""
import argparse
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "molecular_dynamics_b200", "libb200md.so")


def sass_lines(kernel_sub):
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=td, check=True, capture_output=True)
        cubin = [f for f in os.listdir(td) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "--print-line-info", cubin], cwd=td, capture_output=True, text=True).stdout
    out, active, cur = [], False, ("?", 0)
    for line in txt.splitlines():
        if line.startswith(".text."):
            active = kernel_sub in line
            continue
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', line)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            out.append((cur, m.group(2).strip()))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("kernel")
    ap.add_argument("--top", type=int, default=40)
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    body = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
    sass = sass_lines(a.kernel)
    if len(sass) != len(body):
        print(f"warning: {len(sass)} SASS instructions vs {len(body)} ncu rows", file=sys.stderr)
    col = {h: i for i, h in enumerate(hdr)}
    agg = collections.defaultdict(lambda: collections.Counter())
    tot = collections.Counter()
    for (loc, _), r in zip(sass, body):
        d = agg[loc]
        for key, name in (("inst", "Instructions Executed"), ("samples", "# Samples"), ("wave", "L1 Wavefronts Shared"),
                          ("wave_ideal", "L1 Wavefronts Shared Ideal"), ("short_sb", "stall_short_sb"),
                          ("wait", "stall_wait"), ("barrier", "stall_barrier"), ("long_sb", "stall_long_sb"),
                          ("math", "stall_math"), ("branch", "stall_branch_resolving"), ("no_inst", "stall_no_inst")):
            v = float(r[col[name]] or 0)
            d[key] += v
            tot[key] += v
    print(f"total warp instructions {tot['inst']:.3e}, samples {tot['samples']:.0f}, smem wavefronts {tot['wave']:.3e} "
          f"(ideal {tot['wave_ideal']:.3e})")
    print(f"{'source line':32s} {'inst%':>6s} {'smp%':>6s} {'wave%':>6s} {'w/ideal':>7s} {'shortsb':>7s} {'wait':>6s} {'barr':>6s} {'noinst':>6s}")
    for loc, d in sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[: a.top]:
        ratio = d["wave"] / d["wave_ideal"] if d["wave_ideal"] else 0
        print(f"{loc[0] + ':' + str(loc[1]):32s} {100 * d['inst'] / tot['inst']:6.2f} {100 * d['samples'] / tot['samples']:6.2f} "
              f"{100 * d['wave'] / max(tot['wave'], 1):6.2f} {ratio:7.2f} {100 * d['short_sb'] / tot['samples']:7.2f} "
              f"{100 * d['wait'] / tot['samples']:6.2f} {100 * d['barrier'] / tot['samples']:6.2f} {100 * d['no_inst'] / tot['samples']:6.2f}")


if __name__ == "__main__":
    main()
