#!/bin/bash
# ncu --set full capture of the iteration kernel on the default bench workload.  usage: bash scripts/gpu_prof.sh <tag> [workload]
TAG=${1:-rXX}; WL=${2:-cfg2}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python bench.py --workload $WL --steps 50 --warmup 5 --no-cpu-baseline > $OUT/${TAG}_bench_${WL}.json 2> $OUT/${TAG}_bench_${WL}.err; echo "bench rc=$?"
cat $OUT/${TAG}_bench_${WL}.json | cut -c1-400
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sspp_iterate -s 10 -c 1 -f -o $OUT/${TAG}_iterate_full_${WL} \
    python bench.py --workload $WL --steps 20 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_full_${WL}.log 2>&1; echo "ncu full rc=$?"
