#!/bin/bash
# ncu --set full capture of the sweep kernels (one BDF2 step after one warm-up step) on a mid-size workload.
# Usage: bash tools/gpu_ncu.sh <tag> [workload] [ENV=val ...]
TAG=$1; WL=${2:-mid}; shift; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
ARGS="--workload $WL --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
env "$@" timeout 600 python bench.py $ARGS > $OUT/plain.log 2>&1 &&
env "$@" timeout 1500 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 4 -c 4 -o $OUT/sweep_full -f python bench.py $ARGS > $OUT/ncu_full.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/plain.log | cut -c1-600
