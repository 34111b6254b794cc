#!/usr/bin/env python
"""Per-kernel table (markdown) of the metrics the design discussion uses, from an .ncu-rep made
with `ncu --set full`.  Usage: python scripts/ncu_table.py <report.ncu-rep> <out.md> "<title / command>"
Kernels with the same name are listed once per launch in capture order."""
import csv
import subprocess
import sys

rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "smsp__pcsamp_warps_issue_stalled_long_scoreboard", "smsp__pcsamp_warps_issue_stalled_short_scoreboard",
        "smsp__pcsamp_warps_issue_stalled_barrier", "smsp__pcsamp_warps_issue_stalled_wait",
        "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle", "smsp__pcsamp_warps_issue_stalled_lg_throttle",
        "smsp__pcsamp_warps_issue_stalled_mio_throttle", "smsp__pcsamp_warps_issue_stalled_not_selected",
        "smsp__pcsamp_warps_issue_stalled_selected", "smsp__pcsamp_warps_issue_stalled_branch_resolving"]
body = rows[2:]


def short(n):
    n = n.replace("void ", "").replace("vegs::<unnamed>::", "").replace("vegs::", "")
    return n.split("(")[0][-40:]


with open(out, "w") as f:
    f.write(f"# {title}\n\n")
    f.write("| metric | unit | " + " | ".join(f"`{short(r[idx['Kernel Name']])}`" for r in body) + " |\n")
    f.write("|---|---|" + "---|" * len(body) + "\n")
    for w in want:
        if w in idx:
            f.write(f"| {w} | {units[idx[w]]} | " + " | ".join(r[idx[w]] for r in body) + " |\n")
print(open(out).read())
