#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel from an `ncu --set full --import-source on` report.

ncu's CSV source page is per SASS instruction; nvdisasm -g gives the source line of every SASS instruction of the
same cubin (compiled with -lineinfo).  Both list the kernel's instructions in address order, so they are joined by
position.  usage: python tools/ncu_lines.py <rep> <kernel regex> <mangled-name substring> [top N] [launch index]"""
import csv, re, subprocess, sys, tempfile, os, glob

rep, kregex, mangled = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
skip = sys.argv[5] if len(sys.argv) > 5 else "0"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kregex}", "--launch-skip",
                      skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
cols = rows[hdr]
iI, iS = cols.index("Instructions Executed"), cols.index("Source")
iStall = cols.index("Warp Stall Sampling (All Samples)")
sass = [(r[iS].strip(), int(r[iI] or 0), int(r[iStall] or 0)) for r in rows[hdr + 1:] if len(r) > iI and r[0].startswith("0x")]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(root, "gsworld_b200", "csrc", "libgsw.so")], cwd=tmp,
               capture_output=True)
lines = []
for cub in glob.glob(os.path.join(tmp, "*.cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
    m = re.search(r"^(\S*%s\S*):\n" % re.escape(mangled), dis, re.M)
    if not m:
        continue
    cur, started = None, False
    for ln in dis[m.end():].splitlines():
        mm = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if mm:
            cur = (os.path.basename(mm.group(1)), int(mm.group(2)))
            continue
        if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
            lines.append(cur)
            started = True
        elif started and ln.startswith(".L_x") is False and ln.strip().startswith("//-----"):
            break
    break
n = min(len(lines), len(sass))
if abs(len(lines) - len(sass)) > 8:
    print(f"warning: {len(sass)} SASS rows in the report vs {len(lines)} in the cubin (different build?)", file=sys.stderr)
acc, stall = {}, {}
for (src, cnt, st), loc in zip(sass[:n], lines[:n]):
    acc[loc] = acc.get(loc, 0) + cnt
    stall[loc] = stall.get(loc, 0) + st
tot, tst = sum(acc.values()) or 1, sum(stall.values()) or 1
text = {}
print(f"{tot:.4g} warp instructions, {n} SASS instructions; top {top} source lines")
for loc, c in sorted(acc.items(), key=lambda kv: -kv[1])[:top]:
    if loc and loc[0] not in text:
        try:
            text[loc[0]] = open(os.path.join(root, "gsworld_b200", "csrc", loc[0])).read().splitlines()
        except Exception:
            text[loc[0]] = []
    src = text.get(loc[0], [])[loc[1] - 1].strip()[:90] if loc and loc[1] - 1 < len(text.get(loc[0], [])) else ""
    print(f"{100 * c / tot:5.1f}% inst {100 * stall[loc] / tst:5.1f}% stall  {loc[0] if loc else '?'}:{loc[1] if loc else 0:<5d} {src}")
