#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do
for g in on off; do
for pr in -1 0; do
ODVAE_SIDE_PRIORITY=$pr timeout 300 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 --no-roofline --step-graph $g > gpurun_out/abp_${g}_${pr}_${rep}.json 2> gpurun_out/abp.err
python - <<PY
import json
d=json.loads(open('gpurun_out/abp_${g}_${pr}_${rep}.json').read().strip().splitlines()[-1])
print('graph=$g prio=$pr rep=$rep', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['clocks']['sm_mhz'], 'host', round(d['host_enqueue_ms_per_step'],1))
PY
done; done; done
tail -2 gpurun_out/abp.err
