#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_bf16.py tests/test_gpu_timed_path.py tests/test_gpu_edge_cases.py tests/test_gpu_full_size.py -q -x -k "not outcome_rates" > gpurun_out/r02o_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02o_pytest.log; tail -3 gpurun_out/r02o_pytest.log
python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02o.json 2> gpurun_out/r02o.err
python -c "
import json; d=json.loads(open('gpurun_out/r02o.json').read().strip().splitlines()[-1]); print(round(d['value']), round(d['ms_per_step'],1), d['roofline']['frac'], d['roofline']['stage_ms'], d['clocks']['sm_mhz'])"
