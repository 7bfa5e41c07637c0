#!/bin/bash
OUT=gpurun_out/$1; mkdir -p $OUT
T="tests/test_gpu_cell_fields.py tests/test_gpu_temperature.py"
for V in 64 128; do
  RT_CUDA_VAR_THREADS=$V timeout 600 python -m pytest $T -m gpu -x -q > $OUT/pytest_var_$V.log 2>&1; echo "var_threads=$V pytest exit $?"; tail -3 $OUT/pytest_var_$V.log
done
A="--no-cpu-baseline --no-e2e --steps 3 --warmup 3"
for V in 64 128; do
  for W in wide c5r; do
    RT_CUDA_VAR_THREADS=$V timeout 600 python bench.py --workload $W --cell-fields $A > $OUT/bench_cf_${W}_$V.json 2> $OUT/bench_cf_${W}_$V.err; echo "exit $?"
    python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_cf_${W}_$V.json")); print("$W", "$V", "%.3e"%d["value"], d["ms_per_step"], "%.3f"%d["roofline"]["frac"], {k:round(v["GBps"]) for k,v in d["roofline"]["by_kind"].items()}, d["sweep_grid"])
except Exception as e: print("$W $V failed", e)
PY
  done
done
