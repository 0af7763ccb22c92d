#!/usr/bin/env python
"""profiles/ncu_traffic.json from `ncu --set full` reports: per kernel, the mean over the captured
launches of DRAM bytes read + written (what bench.py reports as roofline.traffic) and the
headline metrics, stamped with the commit the capture was taken at.
    python scripts/ncu_traffic.py OUT.json COMMIT "CAPTURE COMMAND" REPORT.ncu-rep [...]"""
import csv
import json
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__waves_per_multiprocessor",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def short(name):
    m = re.search(r"(k_[a-z0-9_]+)(<[^>]*>)?", name)
    base = m.group(1) if m else name
    if base == "k_walk":
        return "k_walk<claim>"
    return base


def main():
    out, commit, cmd, reps = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4:]
    acc = {}
    for rep in reps:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                             text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        hdr, units = rows[0], rows[1]
        col = {h: i for i, h in enumerate(hdr)}
        for r in rows[2:]:
            k = short(r[col["Kernel Name"]])
            e = acc.setdefault(k, {"n": 0, "sum": {}, "units": {}, "kernel": r[col["Kernel Name"]][:90]})
            e["n"] += 1
            for key in KEYS:
                if key in col and r[col[key]] not in ("", "n/a"):
                    e["sum"][key] = e["sum"].get(key, 0.0) + float(r[col[key]].replace(",", ""))
                    e["units"][key] = units[col[key]]
    res = {"commit": commit, "capture": cmd}
    for k, e in acc.items():
        mean = {key: v / e["n"] for key, v in e["sum"].items()}
        rd = mean.get("dram__bytes_read.sum", 0.0) * SCALE.get(e["units"].get("dram__bytes_read.sum", "byte"), 1.0)
        wr = mean.get("dram__bytes_write.sum", 0.0) * SCALE.get(e["units"].get("dram__bytes_write.sum", "byte"), 1.0)
        res[k] = {"dram_bytes_per_launch": rd + wr, "dram_read": rd, "dram_write": wr,
                  "launches_captured": e["n"], "raw": mean, "raw_units": e["units"], "kernel": e["kernel"]}
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    for k, v in res.items():
        if isinstance(v, dict):
            print("%-22s x%d  %8.1f us  dram %7.1f MB" % (k, v["launches_captured"],
                  v["raw"].get("gpu__time_duration.sum", 0), v["dram_bytes_per_launch"] / 1e6))


if __name__ == "__main__":
    main()
