#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -40
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 600 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 600 python bench.py --steps 10 --warmup 3 --step-graph off --extras 0 --cpu-baseline 0 > gpurun_out/bench_eager.json 2> gpurun_out/bench_eager.err; echo "bench eager exit $?"
