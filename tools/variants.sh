#!/bin/bash
# Build variants of libsfpe.so with different compile-time constants (sed on sfpe_fused.cu) into
# gpurun_variants/<name>.so; run them on the GPU box with tools/run_variants.sh.
# usage: tools/variants.sh name 's/^constexpr int ENTRY_BATCH = 3;/constexpr int ENTRY_BATCH = 4;/' [name sed ...]
set -e
cd "$(dirname "$0")/../sensor-fusion-pose-estimation-challenge_b200/csrc"
mkdir -p ../../gpurun_variants
cp sfpe_fused.cu ../../gpurun_variants/.sfpe_fused.orig.cu
while [ $# -ge 2 ]; do
  name=$1; expr=$2; shift 2
  cp ../../gpurun_variants/.sfpe_fused.orig.cu sfpe_fused.cu
  sed -i "$expr" sfpe_fused.cu
  if cmp -s sfpe_fused.cu ../../gpurun_variants/.sfpe_fused.orig.cu && [ "$name" != base ]; then echo "variant $name: sed changed nothing"; fi
  make -j4 > /dev/null 2>&1 || { echo "variant $name: build failed"; cp ../../gpurun_variants/.sfpe_fused.orig.cu sfpe_fused.cu; continue; }
  cp ../libsfpe.so ../../gpurun_variants/$name.so
  echo "variant $name: $(grep -A3 k_fused sfpe_fused.o.ptxas.log | grep -E 'Used' | tr -s ' ')"
done
cp ../../gpurun_variants/.sfpe_fused.orig.cu sfpe_fused.cu
make -j4 > /dev/null 2>&1
