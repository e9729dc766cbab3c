#!/bin/bash
# ncu --set full of one k_train_tc launch at 2^18 samples (after the plain run has exited 0), source page joined with line info
T=${TAG:-r02t}
python scripts/train_once.py 262144 5 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_train_tc" --launch-skip 2 --launch-count 1 -o gpurun_out/prof_${T} python scripts/train_once.py 262144 3 > gpurun_out/${T}_ncu.log 2>&1
ncu -i gpurun_out/prof_${T}.ncu-rep --page raw --csv > gpurun_out/${T}_raw.csv 2>/dev/null
ncu -i gpurun_out/prof_${T}.ncu-rep --page source --csv > gpurun_out/${T}_sass.csv 2>/dev/null
python scripts/ncu_summary.py gpurun_out/${T}_raw.csv | head -40
ls -la gpurun_out | tail -5
