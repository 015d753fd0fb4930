#!/bin/bash
# round-2 GPU call: k_gn_block_pairs timeline, sanitizer passes, config-1 latency, one training iteration on 1 GPU
mkdir -p gpurun_out
python scratch/gn_timing.py > gpurun_out/r02i_gn_timing.log 2>&1; tail -8 gpurun_out/r02i_gn_timing.log
timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python scratch/sanitize_target.py --sims 12 > gpurun_out/r02i_memcheck.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/r02i_memcheck.log; tail -4 gpurun_out/r02i_memcheck.log
timeout 600 compute-sanitizer --tool racecheck --print-limit 20 python scratch/sanitize_target.py --sims 4 --games 8 --no-graph > gpurun_out/r02i_racecheck.log 2>&1
echo "racecheck rc=$?" >> gpurun_out/r02i_racecheck.log; tail -4 gpurun_out/r02i_racecheck.log
timeout 400 compute-sanitizer --tool synccheck --print-limit 20 python scratch/sanitize_target.py --sims 4 --games 8 --no-graph > gpurun_out/r02i_synccheck.log 2>&1
echo "synccheck rc=$?" >> gpurun_out/r02i_synccheck.log; tail -4 gpurun_out/r02i_synccheck.log
python bench.py --mode latency > gpurun_out/r02i_latency.json 2> gpurun_out/r02i_latency.err; echo "latency rc=$?"
python bench.py --mode latency --precision f32 > gpurun_out/r02i_latency_f32.json 2>> gpurun_out/r02i_latency.err
cut -c1-1200 gpurun_out/r02i_latency.json
python bench.py --mode iteration > gpurun_out/r02i_iteration.json 2> gpurun_out/r02i_iteration.err; echo "iteration rc=$?"
cut -c1-1500 gpurun_out/r02i_iteration.json
