#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_timed_path.py tests/test_gpu_edge_cases.py tests/test_gpu_parity.py tests/test_gpu_dropin.py tests/test_gpu_selfplay.py -q -x -k "not outcome_rates" > gpurun_out/r02r_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02r_pytest.log; tail -4 gpurun_out/r02r_pytest.log
for P in 1 0 1 0; do
GZ_PDL=$P timeout 200 python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02r_p$P.json 2> gpurun_out/r02r_p$P.err
python -c "
import json; d=json.loads(open('gpurun_out/r02r_p$P.json').read().strip().splitlines()[-1]); print('pdl $P', round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), d['roofline']['frac'], d['roofline']['stage_ms'], d['clocks']['sm_mhz'])"
done
