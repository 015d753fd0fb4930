#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/r02f_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/r02f_noalloc_alone.log
tail -4 gpurun_out/r02f_noalloc_alone.log
python -m pytest tests -q -m gpu -k "not outcome_rates" > gpurun_out/r02f_pytest_all.log 2>&1
echo "all rc=$?" >> gpurun_out/r02f_pytest_all.log
tail -6 gpurun_out/r02f_pytest_all.log
python __graft_entry__.py --smoke 2>&1 | tail -2
