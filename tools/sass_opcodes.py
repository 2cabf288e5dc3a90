#!/usr/bin/env python
"""Static SASS opcode histogram per kernel of libgsw.so (cuobjdump -sass): what each kernel is made of.
usage: python tools/sass_opcodes.py [gsworld_b200/csrc/libgsw.so] > profiles/rNN_sass_opcodes.txt"""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(root, "gsworld_b200", "csrc", "libgsw.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
cur, hist = None, collections.OrderedDict()
for ln in txt.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = name.replace("void gsw::(anonymous namespace)::", "").replace("gsw::(anonymous namespace)::", "")
        cur = name.split("(gsw::")[0].split("(float")[0].split("(unsigned")[0].split("(int")[0].split("(long")[0].split("(const")[0]
        hist[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", ln)
    if m and cur:
        hist[cur][m.group(1).split(".")[0]] += 1
mark = ("FFMA2", "FMUL2", "FADD2", "UBLKCP", "SYNCS", "MUFU", "ATOMS", "ATOMG", "RED", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "MEMBAR", "CCTL")
for k, c in hist.items():
    tot = sum(c.values())
    top = ", ".join(f"{op} {n}" for op, n in c.most_common(12))
    ev = ", ".join(f"{op} {c[op]}" for op in mark if c.get(op))
    print(f"{k}\n   {tot} instructions; top: {top}\n   of note: {ev}")
