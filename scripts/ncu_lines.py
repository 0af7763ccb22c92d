#!/usr/bin/env python
"""Per-source-line summary of an ncu report (needs -lineinfo and --import-source on):
   python scripts/ncu_lines.py REPORT.ncu-rep [kernel-substring] [top N]
Prints instructions executed and stall samples per CUDA source line, heaviest first, and the
headline raw metrics of every captured launch."""
import csv
import subprocess
import sys

rep = sys.argv[1]
want = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]
if rows:
    hdr = rows[0]
    for k in keys:
        if k in hdr:
            i = hdr.index(k)
            print("%-90s %s (%s)" % (k, [r[i] for r in rows[2:]], rows[1][i]))

out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fpath = fn = None
hdr = None
agg = {}
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        fpath = r[1].split("/")[-1]
        continue
    if len(r) >= 2 and r[0] == "Function Name":
        fn = r[1]
        continue
    if len(r) > 8 and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) == len(hdr) and r[0] != "" and want in (fn or ""):
        ie = hdr.index("Instructions Executed")
        sm = hdr.index("# Samples")
        try:
            agg[(fn, fpath, int(r[0]), r[1].strip()[:90])] = (int(r[ie]), int(r[sm] or 0))
        except ValueError:
            pass
by_fn = {}
for (f, p, l, s), (i, sm) in agg.items():
    by_fn.setdefault(f, []).append((i, sm, p, l, s))
for f, lst in by_fn.items():
    ts = sum(x[1] for x in lst)
    print("== %s: %d samples (inlined lines are counted at every level)" % (f, ts))
    for i, sm, p, l, s in sorted(lst, key=lambda x: -x[1])[:top]:
        print("%9d instr | %5d samples %5.1f%% | %s:%d  %s" % (i, sm, 100.0 * sm / max(ts, 1), p, l, s))
