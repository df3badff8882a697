#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python tools/conv1x1_probe.py 4 > gpurun_out/c1x1_b.log 2>&1; tail -3 gpurun_out/c1x1_b.log
timeout 300 python tools/conv_probe.py 3 > gpurun_out/conv_probe_b.log 2>&1; tail -2 gpurun_out/conv_probe_b.log
PH=64 PC=256 timeout 300 python tools/conv_probe.py 3 > gpurun_out/conv_probe_c.log 2>&1; tail -2 gpurun_out/conv_probe_c.log
timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_kernels_gpu.py -m gpu -x -q > gpurun_out/pytest_engine.log 2>&1; tail -3 gpurun_out/pytest_engine.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; python -c "
import json;d=json.loads(open('gpurun_out/bench_b.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['clocks'])"
