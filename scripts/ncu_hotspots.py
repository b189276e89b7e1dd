#!/usr/bin/env python
"""Aggregate `ncu -i X.ncu-rep --page source --csv --print-source sass,cuda` per CUDA source line.
usage: ncu_hotspots.py src.csv [top_n]"""
import csv, sys
from collections import defaultdict
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None; hdr = None
agg = defaultdict(lambda: [0, 0, 0, ""])   # samples, warp inst, barrier stall
stalls = defaultdict(int)
for row in csv.reader(open(path, errors="replace")):
    if not row: continue
    if row[0] == "File Path": cur_file = row[1].split("/")[-1]; continue
    if row[0] == "Function Name": continue
    if row[0] == "Line No": hdr = row; continue
    if hdr is None or len(row) < len(hdr) - 2: continue
    if row[2] != "-": continue          # SASS row
    try:
        s = int(row[hdr.index("# Samples")]); n = int(row[hdr.index("Instructions Executed")])
    except ValueError: continue
    a = agg[(cur_file, int(row[0]))]; a[0] += s; a[1] += n; a[3] = row[1].strip()[:110]
    for i, h in enumerate(hdr):
        if h.startswith("stall_") and "Not Issued" not in h:
            try: stalls[h] += int(row[i])
            except ValueError: pass
ts = sum(a[0] for a in agg.values()); tn = sum(a[1] for a in agg.values())
print(f"total samples {ts}, warp instructions {tn}")
print("stalls:", ", ".join(f"{k[6:]}={v}" for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:10]))
byfile = defaultdict(lambda: [0, 0])
for (f, l), a in agg.items(): byfile[f][0] += a[0]; byfile[f][1] += a[1]
for f, a in byfile.items(): print(f"  {f}: samples {a[0]} ({100*a[0]/max(ts,1):.1f}%) inst {a[1]} ({100*a[1]/max(tn,1):.1f}%)")
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{f}:{l:4d} samples={a[0]:5d} ({100*a[0]/max(ts,1):4.1f}%) warp_inst={a[1]:9d} ({100*a[1]/max(tn,1):4.1f}%)  {a[3]}")
