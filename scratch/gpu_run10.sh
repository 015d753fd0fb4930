#!/bin/bash
for L in 1 2 3 4; do
python bench.py --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only --lanes $L > gpurun_out/r02k_lanes$L.json 2> gpurun_out/r02k_lanes$L.err
python -c "
import json; d=json.loads(open('gpurun_out/r02k_lanes$L.json').read().strip().splitlines()[-1]); print('lanes $L', round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'])"
done
