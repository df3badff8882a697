#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -30
for i in 1 2; do for t in 0 1; do
ODVAE_PDL=$t timeout 600 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('pdl=$t bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], d['gpu_launches'])"
done; done
