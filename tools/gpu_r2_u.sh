#!/bin/bash
# ncu --set full of the attention score kernels (tc_engine_kernel<4> row statistics / softmax, <5> softmax backward) at the
# bench shape, with and without the stationary B panel
set -x
mkdir -p gpurun_out
timeout 300 python tools/attn_probe.py 64 fp16 || exit 1
for bs in 1 0; do
ODVAE_GEMM_BSTAT=$bs timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tc_engine" -s 6 -c 3 \
    -o gpurun_out/attn_bstat$bs -f python tools/attn_probe.py 64 fp16 > gpurun_out/ncu_attn$bs.log 2>&1
echo "ncu exit $?"
done
ls -la gpurun_out/*.ncu-rep
