#!/bin/bash
# A/B of the cell-dependent tile kernel's shapes: tests (cell fields + temperature update) and the two bench workloads per shape
# Usage: bash tools/gpu_var_ab.sh <tag> ["threads,cpt threads,cpt ..."]
OUT=gpurun_out/$1; mkdir -p $OUT
SHAPES=${2:-"64,4 64,8 128,4"}
T="tests/test_gpu_cell_fields.py tests/test_gpu_temperature.py"
A="--no-cpu-baseline --no-e2e --steps 3 --warmup 3"
for S in $SHAPES; do
  V=${S%,*}; C=${S#*,}
  export RT_CUDA_VAR_THREADS=$V RT_CUDA_VAR_CPT=$C
  timeout 600 python -m pytest $T -m gpu -x -q > $OUT/pytest_var_${V}_$C.log 2>&1; echo "shape $V x $C pytest exit $?"; tail -2 $OUT/pytest_var_${V}_$C.log
  for W in wide c5r; do
    timeout 600 python bench.py --workload $W --cell-fields $A > $OUT/bench_cf_${W}_${V}_$C.json 2> $OUT/bench_cf_${W}_${V}_$C.err; echo "exit $?"
    python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_cf_${W}_${V}_$C.json")); print("$W", "$V x $C", "%.3e"%d["value"], d["ms_per_step"], "%.3f"%d["roofline"]["frac"], {k:round(v["GBps"]) for k,v in d["roofline"]["by_kind"].items()}, d["sweep_grid"])
except Exception as e: print("$W $V x $C failed", e)
PY
  done
done
