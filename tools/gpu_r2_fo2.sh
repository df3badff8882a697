#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_fit_gpu.py tests/test_graph_gpu.py tests/test_patchgan_gpu.py tests/test_kernels_gpu.py tests/test_parity_gpu.py tests/test_lpips_gpu.py -m gpu -x -q > gpurun_out/fo2_pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/fo2_pytest.log
timeout 900 python bench.py --objective full --steps 5 --warmup 3 --extras 0 --cpu-baseline 0 --no-roofline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('full objective', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'host', round(d['host_enqueue_ms_per_step'],1), 'launches/step', d['gpu_launches']/d['steps'])"
timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 --no-roofline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('configs1', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'host', round(d['host_enqueue_ms_per_step'],1), 'launches/step', d['gpu_launches']/d['steps'])"
