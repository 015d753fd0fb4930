#!/bin/bash
# the documented A/B switches still give green parity tests at HEAD
for ENVS in "GZ_GN_PAIRS=0 GZ_GN_FUSED_EMBED=0" "GZ_GN_FUSED_EMBED=0 GZ_GN_SLABS=2 GZ_GN_ALTERNATE=0" "GZ_PDL=1"; do
echo "== $ENVS"
env $ENVS timeout 300 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_timed_path.py tests/test_gpu_edge_cases.py -q -x -k "not outcome_rates and not fp32_bit_exact" 2>&1 | tail -1
done
