#!/usr/bin/env python
"""launches.csv (ncu --metrics gpu__time_duration.sum) -> per-kernel share table."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr, data = rows[h], rows[h + 1:]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in data:
    if len(r) <= vi:
        continue
    name = r[ki].split("(")[0].replace("void ", "").replace("gsw::<unnamed>::", "")[:64]
    v = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1.0)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{len(data)} launches in {steps} steps; {tot / steps:.1f} us of kernel time per step (cold-cache, serialised by ncu)")
print(f"{'us/step':>10} {'launches/step':>14} {'share':>7}  kernel")
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{t / steps:10.1f} {c / steps:14.1f} {100 * t / tot:6.1f}%  {k}")
