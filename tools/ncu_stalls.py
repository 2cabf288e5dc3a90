#!/usr/bin/env python
"""Warp-stall breakdown per kernel from an ncu --set full report (raw page): stalled warps per issued instruction,
by reason, plus occupancy / eligible warps / SM active cycles.  usage: python tools/ncu_stalls.py <rep> [name filter]"""
import csv, subprocess, sys
rep = sys.argv[1]
flt = sys.argv[2] if len(sys.argv) > 2 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
name = hdr.index("Kernel Name")
stall = [(i, h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
         for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled_")]
extra = ["gpu__time_duration.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
         "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_active.avg", "sm__cycles_elapsed.max",
         "launch__waves_per_multiprocessor", "smsp__thread_inst_executed_per_inst_executed.ratio",
         "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps"]
def num(v):
    try:
        return float(v.replace(",", ""))
    except ValueError:
        return 0.0
for r in rows[2:]:
    if flt not in r[name]:
        continue
    print(r[name].split("(")[0].replace("void gsw::<unnamed>::", "")[:60])
    print("   " + ", ".join(f"{h.split('.')[0].replace('smsp__','').replace('launch__','').replace('sm__','')}={r[hdr.index(h)]}" for h in extra if h in hdr))
    top = sorted(((num(r[i]), h) for i, h in stall), reverse=True)[:7]
    print("   stalls/issue: " + ", ".join(f"{h} {v:.2f}" for v, h in top))
