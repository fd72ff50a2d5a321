#!/bin/bash
# GPU box: GPU tests, then bench lines (C2, C3 shapes, phases) for the current build and for every gpurun_variants/*.so
mkdir -p gpurun_out
TAG=${1:-ab}
if [ -z "$SKIP_TESTS" ]; then timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${TAG}_tests.log; fi
shopt -s nullglob
for so in new gpurun_variants/*.so; do
  v=$(basename $so .so)
  if [ "$v" = new ]; then unset SFPE_LIB; else export SFPE_LIB=$PWD/$so; fi
  timeout 300 python bench.py --no-cpu --no-e2e --phases --steps 10 > gpurun_out/${TAG}_bench_$v.json 2> gpurun_out/${TAG}_bench_$v.err; echo "bench $v rc=$?"
  timeout 300 python bench.py --sensor hdl64 --no-cpu --no-e2e --phases --steps 10 > gpurun_out/${TAG}_bench64_$v.json 2> gpurun_out/${TAG}_bench64_$v.err; echo "bench64 $v rc=$?"
done
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_bench*.json')):
    try:
        d=json.load(open(f)); p=d.get('phase_share',{}); t=p.pop('sm_cycles_per_sweep',0)
        print(f.split('/')[-1], '%.0f sweeps/s frac %.3f verified %s'%(d['value'], d['roofline']['frac'], d.get('verified_frames')), 'cyc %.0f'%t, ' '.join('%s %.0f'%(k[:9],v*t) for k,v in p.items()))
    except Exception as e: print(f, 'ERR', e)
PY
