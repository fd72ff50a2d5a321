# C2 + HDL-64E-shaped (phases on) for the current build and every gpurun_variants/*.so
mkdir -p gpurun_out; TAG=${1:-ab2}
shopt -s nullglob
for so in new gpurun_variants/*.so; do
  v=$(basename $so .so); if [ "$v" = new ]; then unset SFPE_LIB; else export SFPE_LIB=$PWD/$so; fi
  timeout 300 python bench.py --no-cpu --no-e2e --no-configs --sustain 0 --phases --steps 10 > gpurun_out/${TAG}_c2_$v.json 2> gpurun_out/${TAG}_c2_$v.err
  timeout 300 python bench.py --config c3 --no-cpu --no-e2e --no-configs --sustain 0 --phases --steps 10 > gpurun_out/${TAG}_c3_$v.json 2> gpurun_out/${TAG}_c3_$v.err
done
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_c*.json')):
    try:
        d=json.load(open(f)); p=d.get('phase_share',{}); t=p.pop('sm_cycles_per_sweep',0)
        print(f.split('/')[-1], '%.0f sweeps/s frac %.3f verified %s'%(d['value'], d['roofline']['frac'], d.get('verified_frames')), 'cyc %.0f'%t, ' '.join('%s %.0f'%(k[:9],v*t) for k,v in p.items()))
    except Exception as e: print(f, 'ERR', e)
PY
