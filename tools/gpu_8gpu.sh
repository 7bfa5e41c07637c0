#!/bin/bash
# 8-GPU box: C5 (default N>1 workload, groups sharded) and the temperature-coupled c5r workload over NCCL
OUT=gpurun_out/$1; mkdir -p $OUT
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541"
timeout 900 $R bench.py --gpus 8 --steps 5 --warmup 3 --no-e2e > $OUT/bench_c5_n8.json 2> $OUT/bench_c5_n8.err; echo "c5 n8 exit $?"
timeout 600 $R bench.py --gpus 8 --workload c5r --cv 1.0 --no-cpu-baseline --no-e2e --steps 3 --warmup 3 > $OUT/bench_cv_c5r_n8.json 2> $OUT/bench_cv_c5r_n8.err; echo "cv n8 exit $?"
timeout 600 $R bench.py --gpus 8 --workload c5r --cell-fields --no-cpu-baseline --no-e2e --steps 3 --warmup 3 > $OUT/bench_cf_c5r_n8.json 2> $OUT/bench_cf_c5r_n8.err; echo "cf n8 exit $?"
python - <<PY
import json
for f in ("bench_c5_n8","bench_cv_c5r_n8","bench_cf_c5r_n8"):
    try:
        d=json.load(open("$OUT/%s.json"%f)); print(f, "%.3e"%d["value"], d["ms_per_step"], d["roofline"]["frac"], d.get("temperature_update"))
    except Exception as e: print(f, "failed", e)
PY
