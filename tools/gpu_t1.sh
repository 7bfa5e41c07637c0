#!/bin/bash
# temperature-update bring-up: the new GPU tests + the cell-field tests (same path), then a short timing of the T-update path
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_temperature.py tests/test_gpu_cell_fields.py -m gpu -x -q --durations=8 > $OUT/pytest_temp.log 2>&1; echo "pytest exit $?"; tail -30 $OUT/pytest_temp.log
