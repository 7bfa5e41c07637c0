#!/bin/bash
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_cell_fields.py tests/test_gpu_temperature.py tests/test_planckian_golden.py tests/test_gpu_parity.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest.log
python tools/planck_bench.py 65536 256 random > $OUT/planck_random.json 2>$OUT/planck.err; cat $OUT/planck_random.json
python tools/planck_bench.py 65536 256 smooth > $OUT/planck_smooth.json 2>>$OUT/planck.err; cat $OUT/planck_smooth.json
for W in wide c5r; do
timeout 600 python bench.py --workload $W --cv 1.0 --no-cpu-baseline --no-e2e --steps 3 --warmup 3 > $OUT/bench_cv_$W.json 2> $OUT/bench_cv_$W.err; echo "exit $?"
python - <<PY
import json
d=json.load(open("$OUT/bench_cv_$W.json")); print("$W %.3e"%d["value"], round(d["ms_per_step"],2), {k:round(v,2) for k,v in d["temperature_update"].items() if k.endswith("_step") and isinstance(v,float)})
PY
done
