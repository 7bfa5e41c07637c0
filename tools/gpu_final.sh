#!/bin/bash
# sanity pass: GPU suite on the current libraries, smoke, then an A/B of library variants on C4 (tools/gpu_ab.sh)
TAG=${1:-final}; shift
mkdir -p gpurun_out/$TAG
timeout 330 python -m pytest tests -m gpu -x -q > gpurun_out/$TAG/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/$TAG/pytest.log
tail -3 gpurun_out/$TAG/pytest.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/$TAG/smoke.log 2>&1; tail -1 gpurun_out/$TAG/smoke.log
[ $# -gt 0 ] && bash tools/gpu_ab.sh $TAG "$@"
