#!/bin/bash
# usage (GPU box): tools/run_variants.sh [bench args]  -- one bench line per gpurun_variants/*.so
for so in gpurun_variants/*.so; do
  SFPE_LIB=$PWD/$so timeout 120 python bench.py --steps 6 --warmup 3 --no-e2e --no-cpu "$@" 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('%-28s %8.0f sweeps/s  %.3f' % ('$so'.split('/')[-1], d['value'], d['roofline']['frac']))"
done
