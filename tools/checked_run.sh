#!/bin/bash
# compute-sanitizer is closed on the GPU pool (the tool answers "closed on this pool").  Stand-in evidence:
# (1) the fused-path tests on a -DSFPE_CHECKED build of the library, where every index the fused kernel
#     computes for a write into shared memory or global scratch is checked against its buffer and traps;
# (2) the poison / guard-cell test (test_fused_rewrites_every_bev_byte) and the repeatability test
#     (test_fused_is_repeatable: a data race between the warps' roles would show as run-to-run differences).
# usage: tools/checked_run.sh build   (build container: makes gpurun_variants/checked.so)
#        tools/checked_run.sh [tag]   (GPU box)
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
  cd sensor-fusion-pose-estimation-challenge_b200/csrc && make -j4 > /dev/null && mkdir -p ../../gpurun_variants &&
  nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-Wall,-fvisibility=hidden -DSFPE_CHECKED -c sfpe_fused.cu -o /tmp/fused_chk.o &&
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../gpurun_variants/checked.so sfpe_api.o sfpe_generic.o /tmp/fused_chk.o sfpe_colormap.o sfpe_post.o sfpe_transport.o sfpe_static.o sfpe_ext.o -lpthread && echo built
  exit $?
fi
TAG=${1:-r02}
mkdir -p gpurun_out
out=gpurun_out/${TAG}_checked_build.txt
echo "== compute-sanitizer on this pool:" > $out
compute-sanitizer --version >> $out 2>&1 | head -3
echo "== fused-path tests on the -DSFPE_CHECKED build (index checks trap):" >> $out
SFPE_LIB=$PWD/gpurun_variants/checked.so timeout 1500 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "fused or bench or host_batch or streaming" >> $out 2>&1
echo "exit code: $?" >> $out
tail -4 $out
