#!/bin/bash
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_warp_kernel -s 1 -c 1 -o $OUT/step_full -f python bench.py --workload wide --no-cpu-baseline --no-e2e --steps 1 --warmup 1 > $OUT/ncu_full.log 2>&1; echo "full exit $?"
ncu -i $OUT/step_full.ncu-rep --page source --csv > $OUT/step_source.csv 2>/dev/null; ls -la $OUT
