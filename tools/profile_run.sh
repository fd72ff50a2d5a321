#!/bin/bash
# ncu evidence for profiles/: launch list of the bench command + one full capture of the fused kernel.
TAG=${1:-r02}
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-configs --sustain 0 > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs --sustain 0 > gpurun_out/${TAG}_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 python bench.py --sweeps 592 --distinct 148 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs --sustain 0 > /dev/null 2>&1; echo "plain592 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused_sweep --launch-skip 3 -c 1 -o gpurun_out/${TAG}_fused -f python bench.py --sweeps 592 --distinct 148 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs --sustain 0 > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
