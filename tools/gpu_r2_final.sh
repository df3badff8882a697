#!/bin/bash
# round-end evidence: tests, smoke, default bench line, ncu launch list of one train step (with DRAM bytes)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; echo "ref exit $?"
export ODVAE_OVERLAP_WGRAD=0
timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --csv --log-file gpurun_out/launches.csv python bench.py --warmup 3 --ncu-step > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
python tools/ncu_pivot.py gpurun_out/launches.csv --dram > gpurun_out/launches_summary.txt 2>&1
head -14 gpurun_out/launches_summary.txt
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ['value','ms_per_step','steps','gpu_launches']}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], d['clocks'])
for k,v in d['extra'].items(): print(k, json.dumps(v)[:400])
print(d['roofline']['hbm_kernels']['groupnorm']); print(d['cpu_baseline'])
print(open('gpurun_out/bench_ref.json').read()[:600])
PY
