#!/bin/bash
mkdir -p gpurun_out
for st in 1 0 1 0; do
echo "== epi_stage=$st (prefetch on)"
ODVAE_EPI_STAGE=$st timeout 300 python tools/conv_probe.py 3 2>&1 | grep "iter 2"
done
echo "== epi_stage=0 prefetch=0"
ODVAE_EPI_STAGE=0 ODVAE_RES_PREFETCH=0 timeout 300 python tools/conv_probe.py 3 2>&1 | grep "iter 2"
for rep in 1 2; do
for st in 1 0; do
ODVAE_EPI_STAGE=$st timeout 300 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 --no-roofline > gpurun_out/abs_${st}_${rep}.json 2> gpurun_out/abs.err
python - <<PY
import json
d=json.loads(open('gpurun_out/abs_${st}_${rep}.json').read().strip().splitlines()[-1])
print('stage=$st rep=$rep', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'])
PY
done; done
