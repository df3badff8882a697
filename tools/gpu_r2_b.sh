#!/bin/bash
# round 2, visit B: PatchGAN + fp16-gradient mode: all GPU tests, gradient error report, a short bench
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_parity_gpu.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -25 gpurun_out/pytest_gpu.log
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q > gpurun_out/pytest_parity.log 2>&1; echo "parity exit $?"
tail -25 gpurun_out/pytest_parity.log
timeout 900 python tests/report_grad_errors.py fp16 > gpurun_out/grad_errors_fp16.txt 2> gpurun_out/grad_errors.err; echo "report exit $?"
grep -v "^    " gpurun_out/grad_errors_fp16.txt
timeout 600 python bench.py --steps 5 --warmup 3 --cpu-baseline 0 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 1500 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
