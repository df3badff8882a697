#!/bin/bash
mkdir -p gpurun_out
for pf in 1 0 1 0; do
echo "== res_prefetch=$pf"
ODVAE_RES_PREFETCH=$pf timeout 300 python tools/conv_probe.py 3 2>&1 | grep "iter 2"
done
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_parity_gpu.py -m gpu -q 2>&1 | tail -3
for rep in 1 2; do
for pf in 1 0; do
ODVAE_RES_PREFETCH=$pf timeout 300 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 --no-roofline > gpurun_out/abr_${pf}_${rep}.json 2> gpurun_out/abr.err
python - <<PY
import json
d=json.loads(open('gpurun_out/abr_${pf}_${rep}.json').read().strip().splitlines()[-1])
print('prefetch=$pf rep=$rep', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'])
PY
done; done
