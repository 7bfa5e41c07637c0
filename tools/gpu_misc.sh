#!/bin/bash
# Planck per-cell kernel timing + ncu fp64 utilisation; bench with e2e phases
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 300 python tools/planck_bench.py 65536 256 > $OUT/planck.json 2> $OUT/planck.err; echo "planck exit $?"; cat $OUT/planck.json
timeout 300 python tools/planck_bench.py 16384 256 > $OUT/planck_small.json 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:planck_integrals -s 1 -c 1 -o $OUT/planck_full -f python tools/planck_bench.py 16384 256 > $OUT/ncu_planck.log 2>&1; echo "ncu exit $?"
timeout 400 python bench.py --steps 5 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('$OUT/bench.json')); print(d['value'], d['e2e'], d['roofline']['frac'])"
