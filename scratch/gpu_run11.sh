#!/bin/bash
ncu --set full --import-source on --clock-control none -k regex:"k_planner_score|k_rollout_tail|k_tree_step|k_planner_pick|k_gnhead_dqn" --launch-skip 200 -c 9 -o gpurun_out/r02l_int \
  python scratch/one_wave.py > gpurun_out/r02l_ncu.log 2>&1
echo rc=$?; ls -la gpurun_out/r02l_int.ncu-rep
