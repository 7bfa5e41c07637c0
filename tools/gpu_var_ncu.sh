#!/bin/bash
OUT=gpurun_out/$1; mkdir -p $OUT
export RT_CUDA_VAR_THREADS=${2:-64}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'var_tile_kernel' -s 4 -c 4 -o $OUT/var_full -f python bench.py --workload wide --cell-fields --no-cpu-baseline --no-e2e --steps 1 --warmup 1 > $OUT/ncu_full.log 2>&1; echo "full exit $?"
ncu -i $OUT/var_full.ncu-rep --page raw --csv > $OUT/var_full_raw.csv 2>/dev/null
python tools/ncu_digest.py $OUT/var_full.ncu-rep > $OUT/var_digest.txt 2>&1; head -60 $OUT/var_digest.txt
