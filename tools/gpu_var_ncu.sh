#!/bin/bash
# ncu --set full of the cell-dependent kernels on the wide workload.  Usage: bash tools/gpu_var_ncu.sh <tag> [kernel regex]
OUT=gpurun_out/$1; mkdir -p $OUT
K=${2:-var_step_kernel}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 2 -o $OUT/var_full -f python bench.py --workload wide --cell-fields --no-cpu-baseline --no-e2e --steps 1 --warmup 1 > $OUT/ncu_full.log 2>&1; echo "full exit $?"
python tools/ncu_digest.py $OUT/var_full.ncu-rep > $OUT/var_digest.txt 2>&1; head -60 $OUT/var_digest.txt | cut -c1-200
