#!/bin/bash
# generic: each argument is "ENV... -- bench args"
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
i=0
for spec in "$@"; do
  i=$((i+1))
  envpart="${spec%%--*}"; argpart="${spec#*--}"
  echo "== [$i] $spec"
  env $envpart timeout 900 python bench.py --no-e2e --no-cpu-baseline $argpart > $OUT/bench_$i.json 2> $OUT/bench_$i.err || tail -5 $OUT/bench_$i.err
  python - "$OUT/bench_$i.json" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    r=d["roofline"]; bk=r["by_kind"]
    print(f"  {d['config']['workload']} value {d['value']:.4e} ms/step {d['ms_per_step']:.3f} frac {r['frac']:.3f} | " + " ".join(f"{k} {v['GBps']:.0f}" for k,v in bk.items()) + f" | grid {d['sweep_grid']} clk {d['clocks']['sm_mhz']} pw {d['clocks'].get('power_w_max')} {d['clocks']['reasons']}")
except Exception as e:
    print("failed:", e)
PY
done
