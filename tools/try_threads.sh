#!/bin/bash
# usage: tools/try_threads.sh "640 768 1024" "0 1"  -- rebuild the library with each CTA size and run the ablation flags
cd sensor-fusion-pose-estimation-challenge_b200/csrc
orig=$(grep "^constexpr int FUSED_THREADS" sfpe_fused.cu)
for T in $1; do
  sed -i "s/^constexpr int FUSED_THREADS = [0-9]*;/constexpr int FUSED_THREADS = $T;/" sfpe_fused.cu
  make -j4 > /dev/null 2>&1 || { echo build failed; exit 1; }
  echo "== FUSED_THREADS $T: $(grep -A2 fused_sweep sfpe_fused.o.ptxas.log | grep -E 'spill' | tr -s ' ')"
  (cd ../..; tools/ablate.sh "$2")
done
sed -i "s/^constexpr int FUSED_THREADS = [0-9]*;/$orig/" sfpe_fused.cu
make -j4 > /dev/null 2>&1
