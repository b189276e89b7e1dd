#!/usr/bin/env python
"""Print the handful of ncu raw-page metrics we track.  usage: ncu_keys.py raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[0], rows[2:]
want = ['gpu__time_duration.sum', 'sm__cycles_active.avg', 'sm__cycles_elapsed.avg', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'idc__requests.sum', 'sm__icc_request_hit_rate.pct',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
for w in want:
    if w in hdr:
        i = hdr.index(w); print(f"{w:85s} {rows[1][i]:12s} {[r[i] for r in data]}")
st = []
for i, h in enumerate(hdr):
    if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio') and 'not_issued' not in h:
        st.append((float(data[0][i]), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]))
print("stalls per issue:", ", ".join(f"{h}={v:.2f}" for v, h in sorted(st, reverse=True)[:9]))
