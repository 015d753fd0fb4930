#!/bin/bash
for V in libgomoku_b200.so libgomoku_b200_nohalves.so libgomoku_b200.so libgomoku_b200_nohalves.so; do
timeout 200 python scratch/ab_lib.py alphazero-gomoku_b200/$V --steps 4 --warmup 3 --selfplay-plies 0 --s400-steps 0 --cpu-seconds 1 --port-only > gpurun_out/r02ab.json 2> gpurun_out/r02ab.err
python -c "
import json; d=json.loads(open('gpurun_out/r02ab.json').read().strip().splitlines()[-1]); print('$V', round(d['value']), round(d['ms_per_step'],1), d['roofline']['stage_ms']['pv_tower'], d['roofline']['stage_ms']['gn_convs'], d['clocks']['sm_mhz'])"
done
