#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_timed_path.py tests/test_gpu_edge_cases.py -q -x -k "not outcome_rates and not fp32_bit_exact" > gpurun_out/r02q_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02q_pytest.log; tail -4 gpurun_out/r02q_pytest.log
timeout 200 python scratch/gn_timing.py 2>&1 | grep gn0 | tail -9
for F in 1 0; do
GZ_GN_FUSED_EMBED=$F timeout 200 python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02q_f$F.json 2> gpurun_out/r02q_f$F.err
python -c "
import json; d=json.loads(open('gpurun_out/r02q_f$F.json').read().strip().splitlines()[-1]); print('fused $F', round(d['value']), round(d['ms_per_step'],1), d['roofline']['frac'], d['roofline']['stage_ms'], d['clocks']['sm_mhz'])"
done
