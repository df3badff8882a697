#!/bin/bash
# visit: GEMM m-tile pairs (attention P / dS GEMMs) -- tests, attention probe A/B, bench A/B
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -x -q -k "gemm" > gpurun_out/mp_pytest.log 2>&1; echo "pytest gemm exit $?"
tail -5 gpurun_out/mp_pytest.log
for m in 1 0 1 0; do
  echo "== attn_probe MPAIR=$m"; ODVAE_GEMM_MPAIR=$m timeout 300 python tools/attn_probe.py 64 fp16 2>&1 | tail -2
done | tee gpurun_out/mp_attn_probe.txt
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/mp_pytest_parity.log 2>&1; echo "pytest parity exit $?"
tail -3 gpurun_out/mp_pytest_parity.log
for round in 1 2; do for m in 0 1; do
  ODVAE_GEMM_MPAIR=$m timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('[mpair=$m] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'roof', round(d['roofline']['frac'],4))"
done; done | tee gpurun_out/mp_bench_ab.txt
