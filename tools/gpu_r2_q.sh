#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -30
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ['value','ms_per_step','steps','gpu_launches','host_enqueue_ms_per_step','cuda_graph']}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], d['clocks'])
for k,v in d['extra'].items(): print(k, json.dumps(v)[:700])
print(d['roofline']['hbm_kernels']['groupnorm']); print(d['cpu_baseline'])
PY
tail -3 gpurun_out/bench.err
