#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 600 python tests/report_bf16_errors.py mixed > gpurun_out/mixed_errors.txt 2>&1; tail -8 gpurun_out/mixed_errors.txt | cut -c1-900
