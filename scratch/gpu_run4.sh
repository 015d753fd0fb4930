#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_bf16.py tests/test_gpu_timed_path.py -q -x -k "not outcome_rates and not fp32_bit_exact" > gpurun_out/r02d_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02d_pytest.log
tail -5 gpurun_out/r02d_pytest.log
python scratch/gn_timing.py > gpurun_out/r02d_gn_timing.log 2>&1; tail -6 gpurun_out/r02d_gn_timing.log
for P in 1 0; do
GZ_GN_PAIRS=$P python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02d_pairs$P.json 2> gpurun_out/r02d_pairs$P.err
python -c "
import json; d=json.loads(open('gpurun_out/r02d_pairs$P.json').read().strip().splitlines()[-1]); print('pairs $P', round(d['value']), round(d['ms_per_step'],1), d['roofline']['frac'], d['roofline']['stage_ms'], d['clocks']['sm_mhz'])"
done
