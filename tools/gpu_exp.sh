#!/bin/bash
# Kernel experiments on the GPU box: bench C4 (device-resident timing only) under several env settings.
# Usage: bash tools/gpu_exp.sh <tag> "ENV1=a ENV2=b" "ENV1=c" ...
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest.log
i=0
for cfg in "$@"; do
  i=$((i+1))
  echo "== [$i] $cfg"
  env $cfg timeout 180 python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu-baseline > $OUT/bench_$i.json 2> $OUT/bench_$i.err || tail -5 $OUT/bench_$i.err
  python - "$OUT/bench_$i.json" "$cfg" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    r=d["roofline"]; bk=r.get("by_kind",{})
    print(f"{sys.argv[2]:44s} value {d['value']:.4e} ms/step {d['ms_per_step']:.2f} frac {r['frac']:.3f} avg_launch_ms {r['avg_launch_ms']:.2f} | " + " ".join(f"{k} {v['GBps']:.0f}" for k,v in bk.items()) + f" | grid {d['sweep_grid']} clk {d['clocks']['sm_mhz']} pw {d['clocks'].get('power_w_max')} {d['clocks']['reasons']}")
except Exception as e:
    print("failed:", e)
PY
done
