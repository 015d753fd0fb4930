#!/bin/bash
# A/B: slab ring depth and alternating board direction in k_gn_block_pairs
mkdir -p gpurun_out
python -m pytest tests/test_gpu_bf16.py tests/test_gpu_timed_path.py -q -x -k "not outcome_rates and not fp32_bit_exact" > gpurun_out/r02m_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02m_pytest.log; tail -3 gpurun_out/r02m_pytest.log
for CFG in "3 1" "2 0" "3 0" "2 1"; do
set -- $CFG
GZ_GN_SLABS=$1 GZ_GN_ALTERNATE=$2 python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02m_s$1a$2.json 2> gpurun_out/r02m_s$1a$2.err
python -c "
import json; d=json.loads(open('gpurun_out/r02m_s$1a$2.json').read().strip().splitlines()[-1]); print('slabs $1 alt $2', round(d['value']), round(d['ms_per_step'],1), d['roofline']['frac'], d['roofline']['stage_ms']['gn_convs'], d['clocks']['sm_mhz'])"
done
