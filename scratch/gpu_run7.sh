#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/r02g_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/r02g_noalloc_alone.log
tail -3 gpurun_out/r02g_noalloc_alone.log
python -m pytest tests/test_gpu_bf16.py tests/test_gpu_edge_cases.py tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/r02g_noalloc_after.log 2>&1
echo "after rc=$?" >> gpurun_out/r02g_noalloc_after.log
tail -3 gpurun_out/r02g_noalloc_after.log
