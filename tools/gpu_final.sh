#!/bin/bash
# sanity pass: GPU suite + smoke on a library (LIBV = variant under csrc/build/ab, default the in-tree library), then an A/B of
# library variants on C4 (tools/gpu_ab.sh).   usage: [LIBV=v7] gpu_final.sh <tag> [variant ...]
TAG=${1:-final}; shift
mkdir -p gpurun_out/$TAG
[ -n "$LIBV" ] && export RT_CUDA_LIB=/root/repo/radiative-transfer_b200/csrc/build/ab/$LIBV.so
timeout 330 python -m pytest tests -m gpu -x -q > gpurun_out/$TAG/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/$TAG/pytest.log
tail -3 gpurun_out/$TAG/pytest.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/$TAG/smoke.log 2>&1; tail -1 gpurun_out/$TAG/smoke.log
unset RT_CUDA_LIB
[ $# -gt 0 ] && bash tools/gpu_ab.sh $TAG "$@"
