#!/bin/bash
# A/B kernel builds of the same ABI: links a variant of ONE csrc file (taken from a git revision) with the current
# objects into radiative-transfer_b200/csrc/build/ab/<tag>.so; select it at run time with RT_CUDA_LIB=<path>.
# Usage: bash tools/ab_build.sh <tag> <git-rev | path of a variant source> <file.cu>
set -e
TAG=$1; REV=$2; F=$3
C=radiative-transfer_b200/csrc; AB=$C/build/ab; mkdir -p $AB
if [ -f "$REV" ]; then cp "$REV" $AB/${TAG}_$F; else git show $REV:$C/$F > $AB/${TAG}_$F; fi
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -ccbin /usr/bin/g++ -Xcompiler -fPIC --fmad=true -I $C -I include $EXTRA"
$NV -c $AB/${TAG}_$F -o $AB/${TAG}_${F%.cu}.o
OBJS=""
for o in rt_capi sweep_kernel step_kernel step_warp_kernel group_terms moments; do
  if [ "$o.cu" == "$F" ]; then OBJS="$OBJS $AB/${TAG}_$o.o"; else OBJS="$OBJS $C/build/$o.o"; fi
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -ccbin /usr/bin/g++ -o $AB/$TAG.so $OBJS -cudart static
echo built $AB/$TAG.so
