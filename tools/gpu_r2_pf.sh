#!/bin/bash
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_kernels_gpu.py -m gpu -q -x 2>&1 | tail -2
for r in 1 2; do for pf in 0 4 8 16; do
ODVAE_GEMM_PREFETCH=$pf timeout 300 python tools/attn_probe.py 64 fp16 | tail -1
done; done
