#!/bin/bash
# Tensor cores vs FP32 FMA for the spline contraction at K = 10 (tools/spline_ab/spline_ab.cu): timings + precision, then ncu pipe shares.
# usage (on the GPU box, repo root): bash scripts/spline_ab.sh <tag>
TAG=${1:-rXX}; OUT=gpurun_out; mkdir -p $OUT
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o tools/spline_ab/spline_ab tools/spline_ab/spline_ab.cu || exit 1
./tools/spline_ab/spline_ab > $OUT/${TAG}_spline_ab.jsonl; echo "run rc=$?"; cat $OUT/${TAG}_spline_ab.jsonl
M=gpu__time_duration.sum,smsp__inst_executed.sum,sm__inst_executed_pipe_tensor.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum
: > $OUT/${TAG}_spline_ab_ncu.csv
# eval_mma<false> launches 221 times before eval_mma<true> does: skip counts pick one warm launch of each
for KS in 'eval_fma 5' 'eval_mma 5' 'eval_mma 230'; do
  set -- $KS
  ncu --clock-control none -k "regex:$1" -s $2 -c 1 --csv --metrics $M ./tools/spline_ab/spline_ab 2>/dev/null | grep -v "^==" | grep -v "^{" >> $OUT/${TAG}_spline_ab_ncu.csv
done
python - <<P
import csv
rows=[r for r in csv.reader(open("$OUT/${TAG}_spline_ab_ncu.csv")) if len(r)>10 and r[0]!="ID"]
for r in rows: print(r[4][:28].ljust(28), r[-3].ljust(58), r[-1])
P
