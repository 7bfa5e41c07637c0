#!/bin/bash
# One GPU-box session: parity tests, smoke, bench (own + reference arm), ncu launch list, ncu full capture of the sweep.
# Usage (from the repo root on the GPU box):  bash tools/gpu_round.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.txt 2>&1
echo "== pytest gpu" ; timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1 ; echo "pytest exit $?" ; tail -5 $OUT/pytest_gpu.log
echo "== smoke" ; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1 ; echo "smoke exit $?" ; tail -2 $OUT/smoke.log
echo "== bench own" ; timeout 400 python bench.py --steps 5 --warmup 3 > $OUT/bench_own.json 2> $OUT/bench_own.err ; echo "bench exit $?" ; cat $OUT/bench_own.json
echo "== bench reference arm" ; timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err ; echo "ref exit $?" ; cat $OUT/bench_ref.json
SMALL="--workload small --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
echo "== ncu launch list (small workload)"
timeout 600 python bench.py $SMALL > $OUT/plain_small.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv python bench.py $SMALL > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
echo "== ncu full capture of the sweep kernels"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 4 -c 4 -o $OUT/sweep_full -f python bench.py $SMALL > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
