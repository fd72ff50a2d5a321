#!/bin/bash
# usage: tools/ablate2.sh "<sweeps list>" "<flags list>" -- phase cycles at different batch sizes
mkdir -p gpurun_out
for n in $1; do for d in $2; do
  python bench.py --sweeps $n --distinct 148 --steps 4 --warmup 3 --no-e2e --no-cpu --phases --debug-flags $d > gpurun_out/dbg2_${n}_$d.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/dbg2_${n}_$d.json')); p=d['phase_share']; t=p.pop('sm_cycles_per_sweep'); print('sweeps %5d flags %3d  %7.0f sweeps/s  total %7.0f  '%($n,$d,d['value'],t)+'  '.join('%s %6.0f'%(k[:9],v*t) for k,v in p.items()))"
done; done
