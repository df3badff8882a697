#!/bin/bash
# round-end evidence (late round 2): tools/gpu_r2_final.sh + ncu --set full of the dominant conv shape (fwd, fwd+residual, dgrad,
# wgrad, fwd+GroupNorm statistics) and of the m-tile-pair attention GEMM P v
bash tools/gpu_r2_final.sh
unset ODVAE_OVERLAP_WGRAD
timeout 300 python tools/conv_probe.py 3 > gpurun_out/conv_probe_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tc_engine" -s 5 -c 5 \
    -o gpurun_out/conv_full -f python tools/conv_probe.py 3 > gpurun_out/ncu_conv.log 2>&1; echo "ncu conv exit $?"
timeout 300 python tools/attn_probe.py 64 fp16 > gpurun_out/attn_probe_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tc_engine" -s 69 -c 2 \
    -o gpurun_out/attn_pv_full -f python tools/attn_probe.py 64 fp16 > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn exit $?"
ls -la gpurun_out/*.ncu-rep
