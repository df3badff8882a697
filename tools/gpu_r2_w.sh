#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -30
timeout 300 python tools/attn_probe.py 64 fp16
for i in 1 2; do for bs in 0 1; do
ODVAE_GEMM_BSTAT=$bs timeout 600 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bstat=$bs', d['value'], d['ms_per_step'], d['clocks']['sm_mhz'])"
done; done
