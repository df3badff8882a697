#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -30
timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 > gpurun_out/bench_ov.json 2> gpurun_out/bench_ov.err; echo "bench exit $?"
ODVAE_OVERLAP_WGRAD=0 timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 > gpurun_out/bench_noov.json 2> gpurun_out/bench_noov.err; echo "bench exit $?"
timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 --step-graph off > gpurun_out/bench_ov_eager.json 2> gpurun_out/bench_ov_eager.err; echo "bench exit $?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_ov.json','gpurun_out/bench_noov.json','gpurun_out/bench_ov_eager.json'):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['engine_ms_per_step'], d['clocks']['sm_mhz'])
    except Exception as e:
        print(f, 'ERR', e)
PY
tail -3 gpurun_out/bench_ov.err
