#!/bin/bash
# temperature-update path: timing on two shapes, launch list, full capture of the Planck integrals + the var sweep
OUT=gpurun_out/$1; mkdir -p $OUT
A="--no-cpu-baseline --no-e2e --steps 3 --warmup 3"
timeout 600 python bench.py --workload wide --cv 1.0 $A > $OUT/bench_cv_wide.json 2> $OUT/bench_cv_wide.err; echo "wide exit $?"
timeout 600 python bench.py --workload c5r --cv 1.0 $A > $OUT/bench_cv_c5r.json 2> $OUT/bench_cv_c5r.err; echo "c5r exit $?"
python - <<PY
import json
for w in ("wide","c5r"):
    try:
        d=json.load(open("$OUT/bench_cv_%s.json"%w)); print(w, "%.3e"%d["value"], d["ms_per_step"], d["roofline"]["frac"], d["temperature_update"])
    except Exception as e: print(w, "failed", e)
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches_cv_wide.csv python bench.py --workload wide --cv 1.0 --no-cpu-baseline --no-e2e --steps 2 --warmup 1 > $OUT/ncu_launch.log 2>&1; echo "launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'planck_integrals|var_sweep|planck_finalize|cell_terms|exchange_partial|moments_kernel' -s 8 -c 9 -o $OUT/cv_full -f python bench.py --workload small --cv 1.0 --no-cpu-baseline --no-e2e --steps 1 --warmup 1 > $OUT/ncu_full.log 2>&1; echo "full exit $?"
ncu -i $OUT/cv_full.ncu-rep --page raw --csv > $OUT/cv_full_raw.csv 2>/dev/null
ls -la $OUT
