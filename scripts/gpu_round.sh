#!/bin/bash
# One GPU-box pass: parity tests, bench lines, ncu launch list, one --set full capture of the hot kernel.
# Usage (from the repo root on the box): bash scripts/gpu_round.sh <tag>
TAG=${1:-rXX}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --workload cfg4 --no-cpu-baseline > $OUT/${TAG}_bench_cfg4.json 2> $OUT/${TAG}_bench_cfg4.err; echo "bench4 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "benchref rc=$?"
cat $OUT/${TAG}_bench.json $OUT/${TAG}_bench_cfg4.json $OUT/${TAG}_bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sspp_iterate -s 10 -c 2 -f -o $OUT/${TAG}_iterate_full \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $OUT
