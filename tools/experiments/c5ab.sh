for so in new gpurun_variants/p192_384.so; do
  v=$(basename $so .so); if [ "$v" = new ]; then unset SFPE_LIB; else export SFPE_LIB=$PWD/$so; fi
  for cfg in c5 c2; do
  timeout 300 python bench.py --config $cfg --no-cpu --no-e2e --no-configs --sustain 0 --steps 20 > gpurun_out/e19_${cfg}_$v.json 2> gpurun_out/e19_${cfg}_$v.err; echo "$cfg $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/e19_${cfg}_$v.json')); print('$cfg $v', round(d['value']), round(d['roofline']['frac'],4), d.get('verified_frames'))"
  done
done
