#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "upsample" > gpurun_out/pytest_up.log 2>&1; echo "up exit $?"
grep -E "^(FAILED|ERROR)|passed|failed|^E  " gpurun_out/pytest_up.log | cut -c1-250 | head -30
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -30
timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 --shapes gpurun_out/shapes.json > gpurun_out/bench_up.json 2> gpurun_out/bench_up.err; echo "bench exit $?"
ODVAE_UPCONV_PHASES=0 timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 > gpurun_out/bench_noup.json 2> gpurun_out/bench_noup.err; echo "bench exit $?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_up.json','gpurun_out/bench_noup.json'):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['engine_ms_per_step'], d['clocks']['sm_mhz'])
    except Exception as e:
        print(f, 'ERR', e)
PY
tail -3 gpurun_out/bench_up.err
