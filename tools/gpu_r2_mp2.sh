#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -x -q -k "gemm" > gpurun_out/mp_pytest.log 2>&1; echo "pytest gemm exit $?"
tail -5 gpurun_out/mp_pytest.log
# attention block at a batch where the pair schedule is on by default vs forced off: same bits expected
ODVAE_GEMM_MPAIR=2 timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" 2>&1 | tail -3
