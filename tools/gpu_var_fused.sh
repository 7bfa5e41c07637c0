#!/bin/bash
# fused vs staged cell-dependent step: tests, then bench of both on the two workloads (+ temperature-coupled)
OUT=gpurun_out/$1; mkdir -p $OUT
T="tests/test_gpu_cell_fields.py tests/test_gpu_temperature.py"
timeout 900 python -m pytest $T -m gpu -x -q > $OUT/pytest_fused.log 2>&1; echo "fused pytest exit $?"; tail -15 $OUT/pytest_fused.log
RT_CUDA_VAR_FUSED=0 timeout 900 python -m pytest $T -m gpu -x -q > $OUT/pytest_staged.log 2>&1; echo "staged pytest exit $?"; tail -3 $OUT/pytest_staged.log
A="--no-cpu-baseline --no-e2e --steps 3 --warmup 3"
for F in 1 0; do
  for W in wide c5r; do
    for X in "--cell-fields" "--cv 1.0"; do
      TAG=$(echo $X | tr -d ' -.'); 
      RT_CUDA_VAR_FUSED=$F timeout 600 python bench.py --workload $W $X $A > $OUT/bench_${TAG}_${W}_f$F.json 2> $OUT/bench_${TAG}_${W}_f$F.err; echo "exit $?"
      python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${TAG}_${W}_f$F.json")); r=d["roofline"]; print("$W", "$X", "fused=$F", "%.3e"%d["value"], round(d["ms_per_step"],2), "%.3f"%r["frac"], round(r["avg_launch_ms"],3), d["sweep_grid"], d.get("temperature_update") and {k:round(v,2) for k,v in d["temperature_update"].items() if k.endswith("_step") and isinstance(v,float)})
except Exception as e: print("$W $X $F failed", e)
PY
    done
  done
done
