#!/bin/bash
# ncu --set full of the round's new kernels: packed-half2 GroupNorm backward (tools/gn_bench.py), one-launch small-map GroupNorm,
# m-tile-pair attention GEMMs (tools/attn_probe.py); each command first runs clean without ncu
mkdir -p gpurun_out
timeout 300 python tools/gn_bench.py 2 > gpurun_out/ncu2_gn_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gn_reduce_h2|gn_bwd_apply_h2|gn_small|gn_apply_kernel" -c 14 \
    -o gpurun_out/gn_h2_full -f python tools/gn_bench.py 1 > gpurun_out/ncu2_gn.log 2>&1; echo "ncu gn exit $?"
timeout 300 python tools/attn_probe.py 64 fp16 > gpurun_out/ncu2_attn_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tc_engine_kernel<3>" -s 8 -c 4 \
    -o gpurun_out/attn_mpair_full -f python tools/attn_probe.py 64 fp16 > gpurun_out/ncu2_attn.log 2>&1; echo "ncu attn exit $?"
ls -la gpurun_out/*.ncu-rep
