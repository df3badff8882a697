#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/attn_probe.py 64 fp16 > gpurun_out/ncu4_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tc_engine" -s 6 -c 2 \
    -o gpurun_out/attn_scores_full -f python tools/attn_probe.py 64 fp16 > gpurun_out/ncu4.log 2>&1; echo "ncu exit $?"
ls -la gpurun_out/attn_scores_full.ncu-rep
