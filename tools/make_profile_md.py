#!/usr/bin/env python
"""profiles/rNN_<kernel>.md from an ncu report: headline counters, stall reasons per issue, region attribution
(tools/ncu_regions.py).  Runs here (no GPU): `ncu -i` only reads the report.

    python tools/make_profile_md.py gpurun_out/r2h_integrate.ncu-rep integrate_kernelILi0ELb0 profiles/r02_integrate_kernel.md "title" ["note"]
"""
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__thread_inst_executed_per_inst_executed.ratio"]


def main():
    rep, kern, out, title = sys.argv[1:5]
    note = sys.argv[5] if len(sys.argv) > 5 else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    lines = [f"# {title}", "", note, "", "| metric | value |", "|---|---|"]
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            lines.append(f"| `{w}` | {vals[i]} {units[i]} |")
    pre, suf = "smsp__average_warps_issue_stalled_", "_per_issue_active.ratio"
    for i, h in enumerate(hdr):
        if h.startswith(pre) and h.endswith(suf) and "not_issued" not in h and float(vals[i]) >= 0.05:
            lines.append(f"| `stall {h[len(pre):-len(suf)]}` | {float(vals[i]):.2f} per issue |")
    reg = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_regions.py"), rep, kern], capture_output=True, text=True).stdout
    keep = [l for l in reg.splitlines() if l and not l.rstrip().endswith("0.00       0.00       0.00")]
    lines += ["", f"Region attribution (`tools/ncu_regions.py {os.path.basename(rep)} {kern}`):", "```"] + keep + ["```", ""]
    with open(out, "w") as fh:
        fh.write("\n".join(lines))
    print("wrote", out)


if __name__ == "__main__":
    main()
