#!/bin/bash
# full GPU pass: every GPU test (outcome-rate tests included), smoke, both bench arms, launch list, --set full of the two GraphNet kernels
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > gpurun_out/r02g_smi.txt
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02g_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02g_pytest.log; tail -3 gpurun_out/r02g_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r02g_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r02g_smoke.log; tail -2 gpurun_out/r02g_smoke.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02g_bench_ref.json 2> gpurun_out/r02g_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02g_bench.json 2> gpurun_out/r02g_bench.err; echo "bench rc=$?"
cut -c1-700 gpurun_out/r02g_bench.json
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 3000 -c 600 --csv --log-file gpurun_out/r02g_launches.csv \
  python bench.py --steps 1 --warmup 1 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02g_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 400 ncu --set full --import-source on --clock-control none -k regex:k_gn_block --launch-skip 40 -c 4 -o gpurun_out/r02g_gn_block \
  python bench.py --steps 1 --warmup 1 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02g_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out | grep r02s
