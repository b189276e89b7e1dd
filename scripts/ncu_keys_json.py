#!/usr/bin/env python
"""ncu raw-page CSV (one or more launches of the iteration kernel) + the explicit FP32 op-count pass -> the record
bench.py attaches to `roofline.ncu` (profiles/r02_iterate_keys.json, one entry per workload).
usage: ncu_keys_json.py <workload> <points_per_launch> <full_raw.csv> <flop_metrics.csv> <out.json>"""
import csv, json, sys
from pathlib import Path

wl, pts, raw, flops, out = sys.argv[1], float(sys.argv[2]), sys.argv[3], sys.argv[4], Path(sys.argv[5])


def table(path):
    rows = list(csv.reader(open(path, errors="replace")))
    i = next(k for k, r in enumerate(rows) if r and r[0] == "ID")
    return rows[i], rows[i + 2:]


def col(hdr, data, name, f=float):
    if name not in hdr:
        return None
    j = hdr.index(name)
    v = [f(r[j].replace(",", "")) for r in data if len(r) > j and r[j] not in ("", "n/a")]
    return sum(v) / len(v) if v else None


hdr, data = table(raw)
rec = dict(source=f"ncu --set full --clock-control none, {len(data)} launch(es) of sspp_iterate_kernel; profiles/" + Path(raw).name,
           issue_active_pct=col(hdr, data, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
           warps_active_pct=col(hdr, data, "sm__warps_active.avg.pct_of_peak_sustained_active"),
           pipe_fma_pct=col(hdr, data, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
           pipe_alu_pct=col(hdr, data, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
           pipe_lsu_pct=col(hdr, data, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
           pipe_xu_pct=col(hdr, data, "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
           warp_inst_per_launch=col(hdr, data, "smsp__inst_executed.sum"),
           lanes_per_inst=col(hdr, data, "smsp__thread_inst_executed_per_inst_executed.ratio"),
           registers=col(hdr, data, "launch__registers_per_thread"),
           duration_us_under_ncu=col(hdr, data, "gpu__time_duration.sum"))
rd, wr = col(hdr, data, "dram__bytes_read.sum"), col(hdr, data, "dram__bytes_write.sum")
j = hdr.index("dram__bytes_read.sum") if "dram__bytes_read.sum" in hdr else None
unit = list(csv.reader(open(raw, errors="replace")))
urow = next(r for k, r in enumerate(unit) if r and r[0] == "ID")
units = unit[unit.index(urow) + 1]
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
if rd is not None:
    rec["dram_bytes_per_launch"] = rd * scale.get(units[hdr.index("dram__bytes_read.sum")], 1) + (wr or 0) * scale.get(units[hdr.index("dram__bytes_write.sum")], 1)
st = []
for i, h in enumerate(hdr):
    if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio") and "not_issued" not in h:
        try:
            st.append((float(data[0][i]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
        except ValueError:
            pass
rec["stalls_per_issue"] = {h: round(v, 2) for v, h in sorted(st, reverse=True)[:6]}
if rec["warp_inst_per_launch"] and rec["lanes_per_inst"]:
    rec["thread_inst_per_point"] = rec["warp_inst_per_launch"] * rec["lanes_per_inst"] / pts
try:                                                   # the --metrics pass logs one row per (launch, metric)
    acc = {}
    rows = list(csv.reader(open(flops, errors="replace")))
    i = next(k for k, r in enumerate(rows) if r and r[0] == "ID")
    hn, hv, hi = rows[i].index("Metric Name"), rows[i].index("Metric Value"), rows[i].index("ID")
    for r in rows[i + 1:]:
        if len(r) > hv:
            acc.setdefault(r[hn], {})[r[hi]] = float(r[hv].replace(",", ""))
    mean = {k: sum(v.values()) / len(v) for k, v in acc.items()}
    fa, fm, ff = (mean.get(f"smsp__sass_thread_inst_executed_op_{o}_pred_on.sum") for o in ("fadd", "fmul", "ffma"))
    if None not in (fa, fm, ff):
        rec.update(fadd_per_point=fa / pts, fmul_per_point=fm / pts, ffma_per_point=ff / pts, flop_per_point_ncu=(fa + fm + 2 * ff) / pts,
                   fp32_share_of_thread_inst=(fa + fm + ff) / mean["smsp__thread_inst_executed.sum"] if "smsp__thread_inst_executed.sum" in mean else None,
                   flop_source="ncu smsp__sass_thread_inst_executed_op_{fadd,fmul,ffma}_pred_on.sum (FMA = 2), whole launch incl. ranking / update")
except (OSError, StopIteration, ValueError):
    pass
allrec = json.loads(out.read_text()) if out.exists() else {}
allrec[wl] = rec
out.write_text(json.dumps(allrec, indent=1) + "\n")
print(json.dumps(rec, indent=1))
