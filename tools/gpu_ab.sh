#!/bin/bash
# A/B bench of library variants built by tools/ab_build.sh (same box, same call): C4, device-resident timing only.
# Usage: bash tools/gpu_ab.sh <tag> <variant> [<variant> ...]     variant = name under csrc/build/ab/ or "cur"
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
for v in "$@"; do
  lib=/root/repo/radiative-transfer_b200/csrc/build/ab/$v.so
  [ "$v" == "cur" ] && lib=/root/repo/radiative-transfer_b200/librt_cuda.so
  RT_CUDA_LIB=$lib timeout 180 python bench.py --workload ${WL:-c4} --steps 3 --warmup 2 --no-e2e --no-cpu-baseline > $OUT/$v.json 2> $OUT/$v.err || tail -3 $OUT/$v.err
  python - $OUT/$v.json $v <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]
    print(f"{sys.argv[2]:10s} value {d['value']:.4e} ms/step {d['ms_per_step']:.2f} frac {r['frac']:.3f} launch_ms {r['avg_launch_ms']:.2f} grid {d['sweep_grid']} pw {d['clocks'].get('power_w_max')}")
except Exception as e: print(sys.argv[2], "failed", e)
PY
done
