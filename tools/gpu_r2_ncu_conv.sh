#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/conv_probe.py 3 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tc_engine" -s 6 -c 6 \
    -o gpurun_out/conv_full -f python tools/conv_probe.py 3 > gpurun_out/ncu_conv.log 2>&1
echo "ncu exit $?"
ls -la gpurun_out/conv_full.ncu-rep
