#!/bin/bash
# Ablation timing of the fused kernel through sfpe_params.reserved[0] debug bits (results are
# invalid with a non-zero flag; this only shows which part of a phase costs the time).  Needs a
# -DSFPE_ABLATE build of the library (SFPE_LIB=...).
# usage: SFPE_LIB=$PWD/gpurun_variants/ablate.so tools/ablate.sh "0 1 8 64" [bench args]
mkdir -p gpurun_out
flags=$1; shift
for d in $flags; do
  python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --phases --debug-flags $d "$@" > gpurun_out/dbg_$d.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/dbg_$d.json')); p=d['phase_share']; t=p.pop('sm_cycles_per_sweep'); print('flags %3d  %7.0f sweeps/s  drop %s total %7.0f  '%($d,d['value'],d.get('dropped_mean'),t)+'  '.join('%s %6.0f'%(k[:9],v*t) for k,v in p.items()))"
done
