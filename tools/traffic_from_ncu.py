#!/usr/bin/env python
"""DRAM read+write bytes per pipeline stage from an `ncu --set full` report of one headline step
-> profiles/<round>_traffic.json (bench.py copies the dominant kernel's figure into roofline.traffic).
usage: python tools/traffic_from_ncu.py gpurun_out/r01_step.ncu-rep profiles/r01_traffic.json"""
import csv, json, subprocess, sys

rep, dst = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, data = rows[0], rows[2:]
iN, iR, iW = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
uR, uW = rows[1][iR], rows[1][iW]
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
stage_of = (("preprocess_kernel<0", "preprocess_static"), ("preprocess_kernel<(bool)0", "preprocess_static"),
            ("preprocess_kernel<1", "preprocess_dynamic"), ("preprocess_kernel<(bool)1", "preprocess_dynamic"),
            ("scan_", "scan"), ("scatter_", "scatter"), ("cull_chunks", "preprocess_static"), ("classify_bins", "sort"), ("refine_big_bins", "sort"),
            ("sort_segments", "sort"), ("count_view_frames", "sort"), ("gather_splats", "gather"),
            ("blend_kernel", "blend"))
acc = {}
for r in data:
    name = r[iN]
    for pat, st in stage_of:
        if pat in name:
            b = float(r[iR].replace(",", "")) * scale[uR] + float(r[iW].replace(",", "")) * scale[uW]
            acc[st] = acc.get(st, 0.0) + b
            break
json.dump({"source": f"{rep} (ncu --set full --clock-control none, one headline step, strict mode; "
                     "tools/traffic_from_ncu.py)", "dram_bytes_per_step": acc}, open(dst, "w"), indent=1)
print(json.dumps(acc, indent=1))
