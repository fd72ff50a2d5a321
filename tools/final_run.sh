#!/bin/bash
# Round-end measurement set on one B200: GPU tests, bench (default config), other configs,
# ncu launch list + one full capture of the fused kernel.  Outputs under gpurun_out/.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/fin_tests.log 2>&1; echo "tests rc=$?"
timeout 600 python bench.py --phases > gpurun_out/fin_bench.json 2> gpurun_out/fin_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --sensor hdl64 --no-cpu --phases > gpurun_out/fin_bench_hdl64.json 2> gpurun_out/fin_bench_hdl64.err; echo "hdl64 rc=$?"
timeout 300 python bench.py --sensor vls128 --bev-res 0.05 --sweeps 64 --distinct 16 --steps 5 --no-cpu --no-e2e --phases > gpurun_out/fin_bench_vls128.json 2> gpurun_out/fin_bench_vls128.err; echo "vls128 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/fin_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/fin_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_fused_sweep --launch-skip 3 -c 1 -o gpurun_out/fin_fused -f python bench.py --sweeps 592 --distinct 148 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/fin_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -2 gpurun_out/fin_tests.log
