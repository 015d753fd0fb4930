#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_edge_cases.py tests/test_gpu_timed_path.py -q -x -k "not outcome_rates and not fp32_bit_exact" > gpurun_out/r02z_pytest.log 2>&1; tail -2 gpurun_out/r02z_pytest.log
timeout 200 python scratch/conv_timing.py --fused 2>&1 | grep -E "pv0" | tail -2
for F in 1 0 1; do
GZ_PV_FUSED_CONV0=$F timeout 200 python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02ab.json 2> gpurun_out/r02ab.err
python -c "
import json; d=json.loads(open('gpurun_out/r02ab.json').read().strip().splitlines()[-1]); st=d['roofline']['stage_ms']; print('pv fused $F', round(d['value']), round(d['ms_per_step'],1), st['pv_tower']+st.get('pv_conv0',0), st['gn_convs'], d['clocks']['sm_mhz'])"
done
