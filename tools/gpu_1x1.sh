#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python tools/conv1x1_probe.py 4 > gpurun_out/c1x1.log 2>&1; cat gpurun_out/c1x1.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_engine -c 2 -o gpurun_out/c1x1_full -f \
    python tools/conv1x1_probe.py 1 > gpurun_out/c1x1_ncu.log 2>&1
ncu -i gpurun_out/c1x1_full.ncu-rep --page raw --csv > gpurun_out/c1x1_raw.csv 2>/dev/null
echo done
