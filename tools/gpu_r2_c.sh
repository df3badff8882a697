#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_parity_gpu.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -40
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q > gpurun_out/pytest_parity.log 2>&1; echo "parity exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_parity.log | head -40
timeout 900 python tests/report_grad_errors.py fp16 > gpurun_out/grad_errors_fp16.txt 2> gpurun_out/grad_errors.err; echo "report exit $?"
grep -v "^    " gpurun_out/grad_errors_fp16.txt
timeout 600 python bench.py --steps 5 --warmup 3 --cpu-baseline 0 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 900 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
