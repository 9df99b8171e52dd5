#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ncu_target_neigh.py && echo "plain ok"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv_q4_0_neigh -s 15 -c 6 -o gpurun_out/r02_neigh_full -f python tools/ncu_target_neigh.py > gpurun_out/r02_ncu_neigh.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r02_neigh_full.ncu-rep --page raw --csv > gpurun_out/r02_neigh_full_raw.csv 2>/dev/null; wc -c gpurun_out/r02_neigh_full_raw.csv
