#!/bin/bash
# ncu --set full of kernels matching a regex during a mid-size bench run (e2e path on so that moments run per step)
# Usage: bash tools/gpu_ncu_k.sh <tag> <kernel-regex> [skip] [count]
OUT=gpurun_out/$1; mkdir -p $OUT
ARGS="--workload ${WL:-mid} --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 300 python bench.py $ARGS > $OUT/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$2 -s ${3:-1} -c ${4:-2} -o $OUT/prof -f python bench.py $ARGS > $OUT/ncu.log 2>&1
echo "ncu exit $?"
