#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none -k regex:"k_" --launch-skip 2155 -c 45 -o /tmp/wave_full python scratch/one_wave.py > gpurun_out/r02v_ncu.log 2>&1
echo "ncu rc=$?"
python profiles/ncu_table.py /tmp/wave_full.ncu-rep > gpurun_out/r02v_wave_table.md
ncu -i /tmp/wave_full.ncu-rep --page raw --csv 2>/dev/null | gzip -9 > gpurun_out/r02v_wave_ncu_raw.csv.gz
ls -la /tmp/wave_full.ncu-rep gpurun_out/r02v_*; cat gpurun_out/r02v_wave_table.md | cut -c1-200
