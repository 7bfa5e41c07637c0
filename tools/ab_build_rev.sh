#!/bin/bash
# Full library build of a git revision for A/B runs: radiative-transfer_b200/csrc/build/ab/<tag>.so
# Usage: bash tools/ab_build_rev.sh <tag> <git-rev>
set -e
TAG=$1; REV=$2
AB=radiative-transfer_b200/csrc/build/ab; D=$AB/${TAG}_tree
rm -rf $D; mkdir -p $D
git archive $REV radiative-transfer_b200/csrc include | tar -x -C $D
make -C $D/radiative-transfer_b200/csrc > /dev/null
cp $D/radiative-transfer_b200/librt_cuda.so $AB/$TAG.so
rm -rf $D
echo built $AB/$TAG.so
