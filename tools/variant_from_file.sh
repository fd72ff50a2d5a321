#!/bin/bash
# Builds gpurun_variants/<name>.so with <file> in place of csrc/sfpe_fused.cu (the tree is left as it was).
# usage: tools/variant_from_file.sh name path/to/alternative_sfpe_fused.cu
set -e
name=$1; alt=$(readlink -f $2)
root=$(cd "$(dirname "$0")/.." && pwd)
csrc=$root/sensor-fusion-pose-estimation-challenge_b200/csrc
mkdir -p $root/gpurun_variants
cp $csrc/sfpe_fused.cu /tmp/.sfpe_fused.keep.cu
cp $alt $csrc/sfpe_fused.cu
if make -C $csrc -j4 > /dev/null 2>&1; then
  cp $csrc/../libsfpe.so $root/gpurun_variants/$name.so
  echo "variant $name: $(grep -A3 k_fused $csrc/sfpe_fused.o.ptxas.log | grep -E 'spill|Used' | tr -s ' ' | tr '\n' ' ')"
else echo "variant $name: build failed"; fi
cp /tmp/.sfpe_fused.keep.cu $csrc/sfpe_fused.cu
make -C $csrc -j4 > /dev/null 2>&1
