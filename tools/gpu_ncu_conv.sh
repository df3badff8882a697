#!/bin/bash
# `ncu --set full` capture of the dominant conv shape (128->128 3x3, batch 64, 256x256): fwd, fwd+residual, dgrad, wgrad
mkdir -p gpurun_out
timeout 300 python tools/conv_probe.py 3 > gpurun_out/conv_probe.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_engine -c 4 -o gpurun_out/conv_full2 -f \
    python tools/conv_probe.py 1 > gpurun_out/ncu_full2.log 2>&1
echo "ncu full exit $?"; cat gpurun_out/conv_probe.log
ncu -i gpurun_out/conv_full2.ncu-rep --page raw --csv > gpurun_out/conv_full2_raw.csv 2>/dev/null
