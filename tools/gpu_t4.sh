#!/bin/bash
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q --durations=5 > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -12 $OUT/pytest_gpu.log
python tools/planck_bench.py 65536 256 random > $OUT/planck_random.json 2>$OUT/planck.err; cat $OUT/planck_random.json
python tools/planck_bench.py 65536 256 smooth > $OUT/planck_smooth.json 2>>$OUT/planck.err; cat $OUT/planck_smooth.json
A="--no-cpu-baseline --no-e2e --steps 3 --warmup 3"
timeout 600 python bench.py --workload wide --cv 1.0 $A > $OUT/bench_cv_wide.json 2> $OUT/bench_cv_wide.err; echo "wide exit $?"
python - <<PY
import json
d=json.load(open("$OUT/bench_cv_wide.json")); print("%.3e"%d["value"], d["ms_per_step"], d["roofline"]["frac"], d["temperature_update"])
PY
