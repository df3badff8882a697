#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_kernels_gpu.py -m gpu -x -q > gpurun_out/pytest_engine.log 2>&1; tail -3 gpurun_out/pytest_engine.log
timeout 300 python tools/conv1x1_probe.py 3 2>&1 | tail -2
timeout 300 python tools/conv_probe.py 3 2>&1 | tail -2
bash tools/gpu_ab_lib.sh
