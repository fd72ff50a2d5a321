#!/bin/bash
# 8-GPU box: the default line (C2, with e2e), C5 and C3 as main lines, one rank per GPU under torchrun.
TAG=${1:-s8}
N=${2:-8}
mkdir -p gpurun_out
run() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err; echo "$name rc=$?"; }
run c2 --no-cpu --no-configs
run c5 --config c5 --no-cpu --no-e2e --no-configs --sustain 0
run c3 --config c3 --no-cpu --no-e2e --no-configs --sustain 0
python - <<PY
import json
for n in ("c2","c5","c3"):
    try:
        d=json.load(open("gpurun_out/${TAG}_%s.json"%n))
        print(n, "value %.0f n_gpus %d frac %.3f verified %s e2e %s clocks %s"%(d["value"], d["n_gpus"], d["roofline"]["frac"], d.get("verified_frames"), (d.get("e2e") or {}).get("by_transport"), d.get("clocks")))
    except Exception as e: print(n, "ERR", e)
PY
