#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_bf16.py -q -x > gpurun_out/r02x_pytest.log 2>&1; tail -2 gpurun_out/r02x_pytest.log
for F in 1 0 1; do
GZ_PV_FUSED_CONV0=$F timeout 200 python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02x_f$F.json 2> gpurun_out/r02x_f$F.err
python -c "
import json; d=json.loads(open('gpurun_out/r02x_f$F.json').read().strip().splitlines()[-1]); print('pv fused $F', round(d['value']), round(d['ms_per_step'],1), d['roofline']['stage_ms']['pv_tower'], d['clocks']['sm_mhz'])"
done
