#!/bin/bash
# Round-end GPU session (r01_v7): full GPU test suite, smoke, bench (own default, reference arm, cell fields, temperature
# update), ncu launch lists of the default command and of the temperature-coupled command, ncu full captures of the
# cell-dependent tile kernel and of the Planck integrals.   Usage: bash tools/gpu_round3.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.txt 2>&1
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_gpu.log
echo "== smoke" ; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1 ; echo "smoke exit $?" ; tail -2 $OUT/smoke.log
echo "== bench own" ; timeout 400 python bench.py --steps 5 --warmup 3 > $OUT/bench_own.json 2> $OUT/bench_own.err ; echo "bench exit $?" ; cut -c1-400 $OUT/bench_own.json
echo "== bench reference arm" ; timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err ; echo "ref exit $?" ; cut -c1-300 $OUT/bench_ref.json
A="--no-cpu-baseline --no-e2e --steps 3 --warmup 3"
for W in wide c5r; do
  timeout 600 python bench.py --workload $W --cell-fields $A > $OUT/bench_cf_$W.json 2> $OUT/bench_cf_$W.err; echo "cf $W exit $?"
  RT_CUDA_VAR_FUSED=0 timeout 600 python bench.py --workload $W --cell-fields $A > $OUT/bench_cf_${W}_staged.json 2> $OUT/bench_cf_${W}_staged.err; echo "cf staged $W exit $?"
  timeout 600 python bench.py --workload $W --cv 1.0 $A > $OUT/bench_cv_$W.json 2> $OUT/bench_cv_$W.err; echo "cv $W exit $?"
done
python - <<PY
import json
for f in ("bench_cf_wide","bench_cf_wide_staged","bench_cf_c5r","bench_cf_c5r_staged","bench_cv_wide","bench_cv_c5r"):
    try:
        d=json.load(open("$OUT/%s.json"%f)); print(f, "%.3e"%d["value"], round(d["ms_per_step"],2), round(d["roofline"]["frac"],3), {k:round(v["GBps"]) for k,v in d["roofline"].get("by_kind",{}).items()}, d.get("temperature_update") and {k:(round(v,2) if isinstance(v,float) else v) for k,v in d["temperature_update"].items() if k.endswith("_step")})
    except Exception as e: print(f, "failed", e)
PY
echo "== ncu launch list (default bench command)"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_default.csv python bench.py --steps 5 --warmup 3 > $OUT/ncu_launches.log 2>&1; echo "ncu launches exit $?"
echo "== ncu launch list (temperature-coupled, wide)"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_cv_wide.csv python bench.py --workload wide --cv 1.0 --no-cpu-baseline --no-e2e --steps 2 --warmup 1 > $OUT/ncu_launches_cv.log 2>&1; echo "ncu launches cv exit $?"
echo "== ncu full: var_tile_kernel (wide, cell fields, stage by stage)"
RT_CUDA_VAR_FUSED=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'var_tile_kernel' -s 4 -c 4 -o $OUT/var_full -f python bench.py --workload wide --cell-fields --no-cpu-baseline --no-e2e --steps 1 --warmup 1 > $OUT/ncu_var.log 2>&1; echo "ncu var exit $?"
python tools/ncu_digest.py $OUT/var_full.ncu-rep > $OUT/var_digest.txt 2>&1
echo "== ncu full: var_step_kernel (wide, cell fields, fused)"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'var_step_kernel' -s 1 -c 2 -o $OUT/var_step_full -f python bench.py --workload wide --cell-fields --no-cpu-baseline --no-e2e --steps 1 --warmup 1 > $OUT/ncu_var_step.log 2>&1; echo "ncu var step exit $?"
python tools/ncu_digest.py $OUT/var_step_full.ncu-rep > $OUT/var_step_digest.txt 2>&1
echo "== ncu full: planck_integrals_kernel + cell_terms (small, cv)"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'planck_integrals|planck_finalize|cell_terms|exchange_partial' -s 4 -c 4 -o $OUT/planck_full -f python bench.py --workload small --cv 1.0 --no-cpu-baseline --no-e2e --steps 1 --warmup 1 > $OUT/ncu_planck.log 2>&1; echo "ncu planck exit $?"
python tools/ncu_digest.py $OUT/planck_full.ncu-rep > $OUT/planck_digest.txt 2>&1
ncu -i $OUT/planck_full.ncu-rep --page raw --csv > $OUT/planck_full_raw.csv 2>/dev/null
rm -f $OUT/var_full.ncu-rep.tmp
ls -la $OUT | head -40
