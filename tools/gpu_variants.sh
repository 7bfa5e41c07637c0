#!/bin/bash
# SURVEY 8(d) variants of the headline run on one GPU: local coupling (dt = 1e-5), no correction source (V = 0),
# the resident variant of c5, and the per-stage path for each.   Usage: bash tools/gpu_variants.sh <tag>
OUT=gpurun_out/$1; mkdir -p $OUT
run() { # name, env, args
  env $2 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline $3 > $OUT/$1.json 2> $OUT/$1.err || tail -3 $OUT/$1.err
  python - $OUT/$1.json $1 <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]; e=d.get("e2e") or {}
    print(f"{sys.argv[2]:22s} value {d['value']:.4e} ms/step {d['ms_per_step']:.2f} frac {r['frac']:.3f} launch_ms {r['avg_launch_ms']:.2f} e2e {e.get('value',0):.3e} batches {d['config']['batches']} clk {d['clocks']['sm_mhz']} {d['clocks']['reasons']}")
except Exception as ex: print(sys.argv[2], "failed", ex)
PY
}
run c4_dt1e-5 "A=1" "--dt 1e-5"
run c4_V0 "A=1" "--V 0"
run c4_dt1e-5_staged "RT_CUDA_FUSED=0" "--dt 1e-5"
run c4_V0_staged "RT_CUDA_FUSED=0" "--V 0"
run c5r "A=1" "--workload c5r"
run c5r_staged "RT_CUDA_FUSED=0" "--workload c5r"
