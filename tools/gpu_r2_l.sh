#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -30
for rep in 1 2; do
for fr in 1 0; do
ODVAE_OVERLAP_FREE=$fr timeout 300 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 --no-roofline > gpurun_out/abf_${fr}_${rep}.json 2> gpurun_out/abf.err
python - <<PY
import json
d=json.loads(open('gpurun_out/abf_${fr}_${rep}.json').read().strip().splitlines()[-1])
print('free=$fr rep=$rep', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['clocks']['sm_mhz'], 'host', round(d['host_enqueue_ms_per_step'],1))
PY
done; done
ODVAE_OVERLAP_FREE=1 timeout 300 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 --no-roofline --step-graph on > gpurun_out/abf_graph.json 2> gpurun_out/abf.err
python - <<PY
import json
d=json.loads(open('gpurun_out/abf_graph.json').read().strip().splitlines()[-1])
print('free=1 graph', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['clocks']['sm_mhz'], 'host', round(d['host_enqueue_ms_per_step'],1))
PY
tail -3 gpurun_out/abf.err
