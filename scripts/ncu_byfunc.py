#!/usr/bin/env python
"""Aggregate an ncu source-page CSV (sass,cuda) by enclosing function of each CUDA source line.
usage: ncu_byfunc.py src.csv"""
import csv, re, sys, bisect
from collections import defaultdict
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent / "sspp_b200" / "csrc"
starts = {}
for f in ("sspp_math.h", "sspp_kernels.cu"):
    lst = []
    for i, line in enumerate(open(ROOT / f), 1):
        m = re.match(r"^(?:template.*\n)?(?:SSPP_HD|__device__|__global__|inline|static)\b[^;(]*?\b(\w+)\s*\(", line)
        if m and not line.startswith(" "): lst.append((i, m.group(1)))
        elif re.match(r"^sspp_iterate_kernel\(", line): lst.append((i, "sspp_iterate_kernel"))
        elif re.match(r"^struct (\w+)", line): lst.append((i, "struct " + re.match(r"^struct (\w+)", line).group(1)))
    starts[f] = lst
def func_of(f, l):
    lst = starts.get(f)
    if not lst: return f
    k = bisect.bisect_right([s for s, _ in lst], l) - 1
    return f"{f}:{lst[k][1]}" if k >= 0 else f
cur = None; hdr = None; agg = defaultdict(lambda: [0, 0])
for row in csv.reader(open(sys.argv[1], errors="replace")):
    if not row: continue
    if row[0] == "File Path": cur = row[1].split("/")[-1]; continue
    if row[0] == "Line No": hdr = row; continue
    if hdr is None or row[0] == "Function Name" or len(row) < 10 or row[2] != "-": continue
    try: s = int(row[hdr.index("# Samples")]); n = int(row[hdr.index("Instructions Executed")])
    except ValueError: continue
    a = agg[func_of(cur, int(row[0]))]; a[0] += s; a[1] += n
ts = sum(a[0] for a in agg.values()); tn = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{k:45s} inst {a[1]:10d} ({100*a[1]/tn:5.1f}%)  samples {a[0]:6d} ({100*a[0]/ts:5.1f}%)")
