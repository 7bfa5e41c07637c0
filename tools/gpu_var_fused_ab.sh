#!/bin/bash
# A/B of the fused cell-dependent step: material-table slots 2 / 1 (tests once, then the two workloads per variant)
OUT=gpurun_out/$1; mkdir -p $OUT
T="tests/test_gpu_cell_fields.py tests/test_gpu_temperature.py"
A="--no-cpu-baseline --no-e2e --steps 3 --warmup 3"
for MB in 2 1; do
  export RT_CUDA_VAR_STEP_MATBUF=$MB
  timeout 900 python -m pytest $T -m gpu -x -q > $OUT/pytest_mb$MB.log 2>&1; echo "matbuf=$MB pytest exit $?"; tail -2 $OUT/pytest_mb$MB.log
  for W in wide c5r; do
    timeout 600 python bench.py --workload $W --cell-fields $A > $OUT/bench_cf_${W}_mb$MB.json 2> $OUT/bench_cf_${W}_mb$MB.err; echo "exit $?"
    python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_cf_${W}_mb$MB.json")); r=d["roofline"]; print("$W", "matbuf=$MB", "%.3e"%d["value"], round(d["ms_per_step"],2), d["sweep_grid"])
except Exception as e: print("$W $MB failed", e)
PY
  done
done
