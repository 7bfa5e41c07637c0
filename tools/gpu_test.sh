#!/bin/bash
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -22 $OUT/pytest_gpu.log
timeout 400 python bench.py --steps 5 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('$OUT/bench.json')); print(d['value'], d['roofline']['frac'], d['e2e'])"
