#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page CSV) into the per-kernel table committed under profiles/."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"),
        ("dram__bytes_write.sum", "dram_wr"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"), ("launch__registers_per_thread", "regs"),
        ("launch__grid_size", "grid"), ("smsp__inst_executed.sum", "warp_inst"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%")]
idx = [(hdr.index(a), b) for a, b in want if a in hdr]
print(" | ".join(f"{b}[{units[i]}]" if units[i] else b for i, b in idx))
for r in data:
    cells = []
    for i, b in idx:
        v = r[i]
        if b == "kernel":
            v = v.split("(")[0].replace("void gsw::<unnamed>::", "")[:34]
        else:
            try:
                v = f"{float(v.replace(',', '')):.4g}"
            except ValueError:
                pass
        cells.append(v)
    print(" | ".join(cells))
