#!/bin/bash
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_cell_fields.py tests/test_gpu_temperature.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest.log
timeout 900 python bench.py --workload wide --cv 1.0 --steps 3 --warmup 3 > $OUT/bench_cv_wide_full.json 2> $OUT/bench_cv_wide_full.err; echo "bench cv full exit $?"; cut -c1-1500 $OUT/bench_cv_wide_full.json; tail -3 $OUT/bench_cv_wide_full.err
