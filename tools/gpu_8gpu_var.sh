#!/bin/bash
# 8-GPU box: 2-rank NCCL temperature test, then the temperature-coupled and the cell-field c5r workloads on 8 ranks
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_temperature_nccl.py -m gpu -x -q > $OUT/pytest_nccl.log 2>&1; echo "nccl pytest exit $?"; tail -2 $OUT/pytest_nccl.log
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541"
A="--gpus 8 --workload c5r --no-cpu-baseline --no-e2e --steps 3 --warmup 3"
timeout 400 $R bench.py $A --cv 1.0 > $OUT/bench_cv_c5r_n8.json 2> $OUT/bench_cv_c5r_n8.err; echo "cv n8 exit $?"
timeout 400 $R bench.py $A --cell-fields > $OUT/bench_cf_c5r_n8.json 2> $OUT/bench_cf_c5r_n8.err; echo "cf n8 exit $?"
python - <<PY
import json
for f in ("bench_cv_c5r_n8","bench_cf_c5r_n8"):
    try:
        d=json.load(open("$OUT/%s.json"%f)); print(f, "%.3e"%d["value"], round(d["ms_per_step"],2), d.get("temperature_update") and {k:round(v,2) for k,v in d["temperature_update"].items() if k.endswith("_step") and isinstance(v,float)})
    except Exception as e: print(f, "failed", e)
PY
