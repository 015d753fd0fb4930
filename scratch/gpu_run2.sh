#!/bin/bash
mkdir -p gpurun_out
python scratch/gn_timing.py > gpurun_out/r02b_gn_timing.log 2>&1
python -m pytest tests/test_gpu_bf16.py tests/test_gpu_arena_update.py tests/test_gpu_parity.py tests/test_gpu_selfplay.py tests/test_gpu_dropin.py -q -x > gpurun_out/r02b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02b_pytest.log
python bench.py --steps 5 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err
echo "bench rc=$?" >> gpurun_out/r02b_bench.err
tail -3 gpurun_out/r02b_pytest.log; tail -8 gpurun_out/r02b_gn_timing.log
python -c "
import json; d=json.loads(open('gpurun_out/r02b_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['stage_ms'], d['clocks'])"
