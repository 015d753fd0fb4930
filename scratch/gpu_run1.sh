#!/bin/bash
# round-2 GPU call 1: full GPU test suite, smoke, bench (engine + reference arm + latency), k_gn_block timeline, sanitizers
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02a_smi.txt
nproc >> gpurun_out/r02a_smi.txt
python -m pytest tests -m gpu -q -k "not outcome_rates" > gpurun_out/r02a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02a_pytest.log
python __graft_entry__.py --smoke > gpurun_out/r02a_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r02a_smoke.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err
echo "bench rc=$?" >> gpurun_out/r02a_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02a_bench_ref.json 2> gpurun_out/r02a_bench_ref.err
python bench.py --mode latency > gpurun_out/r02a_latency.json 2> gpurun_out/r02a_latency.err
python bench.py --mode latency --precision f32 > gpurun_out/r02a_latency_f32.json 2>> gpurun_out/r02a_latency.err
python scratch/gn_timing.py > gpurun_out/r02a_gn_timing.log 2>&1
timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python scratch/sanitize_target.py --sims 12 > gpurun_out/r02a_memcheck.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/r02a_memcheck.log
timeout 420 compute-sanitizer --tool racecheck --print-limit 20 python scratch/sanitize_target.py --sims 4 --games 8 > gpurun_out/r02a_racecheck.log 2>&1
echo "racecheck rc=$?" >> gpurun_out/r02a_racecheck.log
ls -la gpurun_out/test_reports 2>/dev/null
tail -3 gpurun_out/r02a_pytest.log
