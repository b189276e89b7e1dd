#!/bin/bash
# One GPU-box pass: parity tests, bench lines, ncu launch list, --set full capture + FP32 op counts of the hot kernel.
# Usage (from the repo root on the box): bash scripts/gpu_round.sh <tag> [quick]
TAG=${1:-rXX}; MODE=${2:-full}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 600 python bench.py > $OUT/${TAG}_bench_cfg2.json 2> $OUT/${TAG}_bench_cfg2.err; echo "bench rc=$?"
cut -c1-1500 $OUT/${TAG}_bench_cfg2.json
[ "$MODE" = quick ] && exit 0
for WL in cfg1 cfg4 cfg3 cfg5; do
  timeout 900 python bench.py --workload $WL --steps 50 --warmup 5 > $OUT/${TAG}_bench_${WL}.json 2> $OUT/${TAG}_bench_${WL}.err; echo "bench $WL rc=$?"
  cut -c1-600 $OUT/${TAG}_bench_${WL}.json
done
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/${TAG}_bench_reference_arm.json 2> $OUT/${TAG}_bench_ref.err; echo "benchref rc=$?"
cut -c1-400 $OUT/${TAG}_bench_reference_arm.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sspp_iterate -s 10 -c 2 -f -o /tmp/${TAG}_iterate_full \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/${TAG}_iterate_full.ncu-rep --page raw --csv > $OUT/${TAG}_iterate_full_raw.csv 2>/dev/null
timeout 600 ncu --metrics smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__thread_inst_executed.sum,smsp__inst_executed.sum \
    --clock-control none -k regex:sspp_iterate -s 10 -c 2 --csv --log-file $OUT/${TAG}_iterate_flops.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_flops.log 2>&1; echo "ncu flops rc=$?"
python scripts/ncu_keys_json.py cfg2 262144 $OUT/${TAG}_iterate_full_raw.csv $OUT/${TAG}_iterate_flops.csv $OUT/${TAG}_iterate_keys.json
python scripts/ncu_keys.py $OUT/${TAG}_iterate_full_raw.csv > $OUT/${TAG}_iterate_summary.txt 2>&1
ncu -i /tmp/${TAG}_iterate_full.ncu-rep --page source --csv --print-source sass,cuda > /tmp/${TAG}_iterate_source.csv 2>/dev/null
python scripts/ncu_byfunc.py /tmp/${TAG}_iterate_source.csv >> $OUT/${TAG}_iterate_summary.txt 2>&1
python scripts/phase_times.py > $OUT/${TAG}_phase_timeline.json 2>$OUT/${TAG}_phase.err
# the .ncu-rep and the source page (35 MB each) stay on the box: gpurun_out/ is capped at 64 MiB
ls -la $OUT | tail -30
