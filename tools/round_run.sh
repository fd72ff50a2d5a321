#!/bin/bash
# Round measurement set on one B200: GPU tests, default bench line (with configs), reference arm, launch list.
TAG=${1:-r02}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${TAG}_tests.log
timeout 900 python bench.py --phases > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err; echo "ref rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/${TAG}_bench.json'))
print('value %.0f frac %.3f verified %s sustained %s'%(d['value'],d['roofline']['frac'],d.get('verified_frames'),d.get('sustained')))
print('e2e', d.get('e2e',{}).get('by_transport'))
print('cpu', {k:v for k,v in d.get('cpu_baseline',{}).items() if k!='sample'})
for k,v in d.get('configs',{}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a not in ('note','entry')})
r=json.load(open('gpurun_out/${TAG}_ref.json')); print('ref', r['value'], r['cpu_baseline']['kind'], r['cpu_baseline']['cores'])
PY
