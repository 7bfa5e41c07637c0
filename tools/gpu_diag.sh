#!/bin/bash
# diagnostic pass: the fused-step tests on library variants under csrc/build/ab (RT_CUDA_LIB), "cur" = the in-tree library
# usage: gpu_diag.sh <tag> <variant>[:ENV=VALUE] ...
OUT=gpurun_out/$1; shift; mkdir -p $OUT
AB=/root/repo/radiative-transfer_b200/csrc/build/ab
for spec in "$@"; do
  v=${spec%%:*}; envs=""; [ "$spec" != "$v" ] && envs=${spec#*:}
  lib=$AB/$v.so; [ "$v" == "cur" ] && lib=/root/repo/radiative-transfer_b200/librt_cuda.so
  env RT_CUDA_LIB=$lib $envs timeout -s KILL 40 python -m pytest tests/test_gpu_fused.py -x -q > $OUT/fused_${spec//[:=]/_}.log 2>&1
  echo "$spec rc=$? $(tail -1 $OUT/fused_${spec//[:=]/_}.log | cut -c1-150)"
done
