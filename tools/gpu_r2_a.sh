#!/bin/bash
# round 2, visit A: per-tensor gradient difference errors (all modes) + parity tests with the new fp32 1e-4 bound
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python tests/report_grad_errors.py fp32 mixed bf16 > gpurun_out/grad_errors.txt 2> gpurun_out/grad_errors.err; echo "report exit $?"
tail -30 gpurun_out/grad_errors.txt
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q > gpurun_out/pytest_parity.log 2>&1; echo "pytest exit $?"
tail -15 gpurun_out/pytest_parity.log
