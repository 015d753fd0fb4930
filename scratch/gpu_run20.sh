#!/bin/bash
# N=8: BASELINE config 4 (8 x 8192 = 65 536 concurrent games: sims/s and steady-state self-play games/h) and config 5 (one training iteration)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29521 bench.py --gpus 8 --games 8192 --steps 2 --warmup 3 --s400-steps 0 --selfplay-plies 80 > gpurun_out/r02u_config4_8gpu.json 2> gpurun_out/r02u_config4_8gpu.err
echo "config4 rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02u_config4_8gpu.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['selfplay_games_per_hour'], d['clocks'])"
timeout 420 $TR --master-port 29522 bench.py --gpus 8 --mode iteration --games 4096 --min-games 4096 --eval-games 64 > gpurun_out/r02u_iteration_8gpu.json 2> gpurun_out/r02u_iteration_8gpu.err
echo "iteration rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02u_iteration_8gpu.json').read().strip().splitlines()[-1]); print(d['phases_s'], d['games'], d['value'], d['clocks'])"
tail -2 gpurun_out/r02u_iteration_8gpu.err
