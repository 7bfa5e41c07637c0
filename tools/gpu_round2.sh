#!/bin/bash
# One GPU-box session for the fused-step build: smoke, bench (own + reference arm), ncu launch list of the default bench
# command, ncu full capture of the fused step kernel on the "wide" workload (as many rows as a resident c4 batch).
# Usage (from the repo root on the GPU box):  bash tools/gpu_round2.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.txt 2>&1
echo "== smoke" ; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1 ; echo "smoke exit $?" ; tail -2 $OUT/smoke.log
echo "== bench own" ; timeout 400 python bench.py --steps 5 --warmup 3 > $OUT/bench_own.json 2> $OUT/bench_own.err ; echo "bench exit $?" ; cat $OUT/bench_own.json
echo "== bench own, per-stage kernels (RT_CUDA_FUSED=0)" ; RT_CUDA_FUSED=0 timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_staged.json 2> $OUT/bench_staged.err ; echo "bench exit $?" ; cat $OUT/bench_staged.json
echo "== bench reference arm" ; timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err ; echo "ref exit $?" ; cat $OUT/bench_ref.json
echo "== ncu launch list (default bench command)"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv python bench.py --steps 5 --warmup 3 > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
echo "== ncu full capture of the fused step kernel (wide workload)"
WIDE="--workload wide --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 300 python bench.py $WIDE > $OUT/plain_wide.json 2> $OUT/plain_wide.err &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:step_warp_kernel -s 1 -c 2 -o $OUT/step_full -f python bench.py $WIDE > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
