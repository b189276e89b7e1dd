#!/bin/bash
# Multi-GPU pass on one box: usage bash scripts/gpu_round_multi.sh <tag> <N>
TAG=${1:-rXX}; N=${2:-8}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29541 bench.py --gpus $N --steps 100 --warmup 10 > $OUT/${TAG}_bench_cfg2_${N}gpu.json 2> $OUT/${TAG}_bench_cfg2_${N}gpu.err; echo "cfg2 rc=$?"
timeout 600 $TR --master-port 29542 bench.py --gpus $N --workload cfg3 --steps 50 --warmup 5 > $OUT/${TAG}_bench_cfg3_${N}gpu.json 2> $OUT/${TAG}_bench_cfg3_${N}gpu.err; echo "cfg3 rc=$?"
timeout 600 $TR --master-port 29543 bench.py --gpus $N --workload cfg5 --steps 5 --warmup 1 > $OUT/${TAG}_bench_cfg5_${N}gpu.json 2> $OUT/${TAG}_bench_cfg5_${N}gpu.err; echo "cfg5 rc=$?"
timeout 600 $TR --master-port 29544 bench.py --gpus $N --workload cfg4 --steps 20 --warmup 3 > $OUT/${TAG}_bench_cfg4_${N}gpu.json 2> $OUT/${TAG}_bench_cfg4_${N}gpu.err; echo "cfg4 rc=$?"
POP_PER_RANK=1024 POP_SOAK=${SOAK:-100000} timeout 900 $TR --master-port 29545 scripts/check_popshard.py 2>&1 | grep -E "OK|FAIL|phases|unsharded|soak" > $OUT/${TAG}_popshard_soak_${N}gpu.txt; echo "soak rc=$?"
POP_CASE=sdf timeout 600 $TR --master-port 29546 scripts/check_popshard.py 2>&1 | grep -E "OK|FAIL|phases|unsharded" > $OUT/${TAG}_cfg3_popshard_${N}gpu.txt
cat $OUT/${TAG}_popshard_soak_${N}gpu.txt $OUT/${TAG}_cfg3_popshard_${N}gpu.txt
for f in cfg2 cfg3 cfg5 cfg4; do cut -c1-260 $OUT/${TAG}_bench_${f}_${N}gpu.json; done
