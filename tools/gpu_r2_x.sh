#!/bin/bash
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python tools/attn_probe.py 64 fp16
for i in 1 2; do
timeout 600 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', d['value'], d['ms_per_step'], d['clocks']['sm_mhz'])"
done
