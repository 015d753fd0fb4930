#!/bin/bash
# N=2: one training iteration (config 5, data-parallel update) and the search line
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --mode iteration --games 4096 --min-games 4096 --eval-games 64 > gpurun_out/r02t_iteration_2gpu.json 2> gpurun_out/r02t_iteration_2gpu.err
echo "iteration rc=$?"; cut -c1-300 gpurun_out/r02t_iteration_2gpu.json; python -c "
import json; d=json.loads(open('gpurun_out/r02t_iteration_2gpu.json').read().strip().splitlines()[-1]); print(d['phases_s'], d['games'], d['value'])"
tail -3 gpurun_out/r02t_iteration_2gpu.err
