#!/bin/bash
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:var_step_kernel -s 1 -c 1 -o $OUT/var_step_full -f python bench.py --workload wide --cell-fields --no-cpu-baseline --no-e2e --steps 1 --warmup 1 > $OUT/ncu_full.log 2>&1; echo "full exit $?"
python tools/ncu_digest.py $OUT/var_step_full.ncu-rep 2>&1 | head -24 | cut -c1-170
