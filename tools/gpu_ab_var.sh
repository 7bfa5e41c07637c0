#!/bin/bash
# A/B bench of library variants (tools/ab_build.sh style .so files) on the cell-dependent workloads
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
for v in "$@"; do
  lib=/root/repo/radiative-transfer_b200/csrc/build/ab/$v.so
  [ "$v" == "cur" ] && lib=/root/repo/radiative-transfer_b200/librt_cuda.so
  for W in wide c5r; do
    RT_CUDA_LIB=$lib timeout 300 python bench.py --workload $W --cell-fields --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${v}_$W.json 2> $OUT/${v}_$W.err || tail -3 $OUT/${v}_$W.err
    python - $OUT/${v}_$W.json $v $W <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]
    print(f"{sys.argv[2]:10s} {sys.argv[3]:6s} value {d['value']:.4e} ms/step {d['ms_per_step']:.2f} launch_ms {r['avg_launch_ms']:.2f} grid {d['sweep_grid']}")
except Exception as e: print(sys.argv[2], "failed", e)
PY
  done
done
