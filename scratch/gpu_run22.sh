#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_edge_cases.py tests/test_gpu_timed_path.py tests/test_gpu_parity.py -q -x -k "not outcome_rates and not fp32_bit_exact" > gpurun_out/r02w_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02w_pytest.log; tail -12 gpurun_out/r02w_pytest.log
for F in 1 0 1 0; do
GZ_PV_FUSED_CONV0=$F timeout 200 python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02w_f$F.json 2> gpurun_out/r02w_f$F.err
python -c "
import json; d=json.loads(open('gpurun_out/r02w_f$F.json').read().strip().splitlines()[-1]); print('pv fused $F', round(d['value']), round(d['ms_per_step'],1), d['roofline']['frac'], d['roofline']['stage_ms'], d['clocks']['sm_mhz'])"
done
