#!/bin/bash
# parity tests + bench line + per-shape engine table (no ncu)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 --cpu-baseline 0 --shapes gpurun_out/shapes.json > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 1200 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
