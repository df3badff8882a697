#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -30
timeout 300 python tools/gn_bench.py 5 2>&1 | tail -4
for rep in 1 2; do
timeout 300 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 > gpurun_out/bench_gn_$rep.json 2> gpurun_out/bench_gn.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_gn_$rep.json').read().strip().splitlines()[-1])
print('rep=$rep', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['clocks']['sm_mhz'], 'host', round(d['host_enqueue_ms_per_step'],1), 'roof', round(d['roofline']['frac'],3), {k:(round(v['ms'],2), round(v['frac_of_hbm_peak'],3)) for k,v in d['roofline']['hbm_kernels']['groupnorm'].items()})
PY
done
tail -3 gpurun_out/bench_gn.err
