#!/bin/bash
# compute-sanitizer passes over the CUDA kernels ("race detection"): memcheck,
# racecheck (shared-memory hazards: the fused kernel's inter-warp protocol over flags in shared memory
# and named barrier 1) and synccheck, on small-frame subsets of the GPU parity tests.
# usage (GPU box): tools/sanitize.sh [tag]   -> gpurun_out/<tag>_sanitizer_{memcheck,racecheck,synccheck}.txt
TAG=${1:-r02}
mkdir -p gpurun_out
SEL='fused_path_ragged or fused_path_invalid or fused_path_saturating or fused_rewrites_every_bev_byte or fused_path_vs_reference_golden or fused_path_adversarial'
for tool in memcheck racecheck synccheck; do
  out=gpurun_out/${TAG}_sanitizer_${tool}.txt
  echo "== compute-sanitizer --tool $tool: tests/test_gpu_parity.py -k '$SEL' + tests/test_gpu_post.py" > $out
  timeout 1500 compute-sanitizer --tool $tool --print-limit 30 --error-exitcode 86 \
    python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "$SEL" >> $out 2>&1
  echo "exit code (parity subset): $?" >> $out
  if [ $tool != racecheck ]; then
    timeout 900 compute-sanitizer --tool $tool --print-limit 30 --error-exitcode 86 \
      python -m pytest tests/test_gpu_post.py -x -q -m gpu >> $out 2>&1
    echo "exit code (post kernels): $?" >> $out
  fi
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|exit code|passed|failed" $out | tail -8
done
