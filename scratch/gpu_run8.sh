#!/bin/bash
# round-2 GPU call (re-entry): GPU tests, smoke, bench both arms, launch list, --set full of the GraphNet block kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02h_smi.txt
nproc >> gpurun_out/r02h_smi.txt
python -m pytest tests -m gpu -q -x -k "not outcome_rates" > gpurun_out/r02h_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02h_pytest.log
tail -3 gpurun_out/r02h_pytest.log
python __graft_entry__.py --smoke > gpurun_out/r02h_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r02h_smoke.log
tail -2 gpurun_out/r02h_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02h_bench_ref.json 2> gpurun_out/r02h_bench_ref.err
echo "ref rc=$?"
python bench.py --steps 5 --warmup 3 > gpurun_out/r02h_bench.json 2> gpurun_out/r02h_bench.err
echo "bench rc=$?"
cat gpurun_out/r02h_bench.json | cut -c1-1500
for P in 0; do
GZ_GN_PAIRS=$P python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02h_pairs$P.json 2> gpurun_out/r02h_pairs$P.err
done
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 3000 -c 600 --csv --log-file gpurun_out/r02h_launches.csv \
  python bench.py --steps 1 --warmup 1 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02h_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --import-source on --clock-control none -k regex:k_gn_block --launch-skip 40 -c 2 -o gpurun_out/r02h_gn_block \
  python bench.py --steps 1 --warmup 1 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02h_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
