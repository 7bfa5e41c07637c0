#!/bin/bash
# 2-GPU box: NCCL temperature test, then the temperature-coupled c5r workload at N=1 and N=2 (strong scaling of the extension)
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 500 python -m pytest tests/test_gpu_temperature_nccl.py -m gpu -x -q > $OUT/pytest_nccl.log 2>&1; echo "nccl pytest exit $?"; tail -3 $OUT/pytest_nccl.log
A="--workload c5r --cv 1.0 --no-cpu-baseline --no-e2e --steps 3 --warmup 3"
timeout 600 python bench.py --gpus 1 $A > $OUT/bench_cv_c5r_n1.json 2> $OUT/bench_cv_c5r_n1.err; echo "n1 exit $?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 $A > $OUT/bench_cv_c5r_n2.json 2> $OUT/bench_cv_c5r_n2.err; echo "n2 exit $?"
python - <<PY
import json
for n in (1,2):
    try:
        d=json.load(open("$OUT/bench_cv_c5r_n%d.json"%n)); print(n, "%.3e"%d["value"], d["ms_per_step"], d["temperature_update"])
    except Exception as e: print(n, "failed", e)
PY
tail -5 $OUT/bench_cv_c5r_n2.err
