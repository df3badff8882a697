#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "attention" 2>&1 | tail -5
ODVAE_GEMM_BSTAT=0 timeout 300 python tools/attn_probe.py 64 fp16
ODVAE_GEMM_BSTAT=1 timeout 300 python tools/attn_probe.py 64 fp16
ODVAE_GEMM_BSTAT=0 timeout 300 python tools/attn_probe.py 64 fp16
ODVAE_GEMM_BSTAT=1 timeout 300 python tools/attn_probe.py 64 fp16
