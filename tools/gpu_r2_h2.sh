#!/bin/bash
# visit: packed-half2 GroupNorm backward -- kernel tests, per-kernel A/B of the build variants, parity tests, bench A/B
mkdir -p gpurun_out
V=/root/repo/tools/experiments/variants
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k groupnorm > gpurun_out/h2_pytest_gn.log 2>&1; echo "pytest gn exit $?"
tail -3 gpurun_out/h2_pytest_gn.log
for n in default base h2b h2c default base; do
  if [ $n = default ]; then unset ODVAE_LIB_PATH; else export ODVAE_LIB_PATH=$V/lib_gn_$n.so; fi
  echo "== gn_bench $n"; timeout 300 python tools/gn_bench.py 5 2>&1 | tail -4
done | tee gpurun_out/h2_gn_bench.txt
unset ODVAE_LIB_PATH
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_patchgan_gpu.py tests/test_graph_gpu.py -m gpu -x -q > gpurun_out/h2_pytest_parity.log 2>&1; echo "pytest parity exit $?"
tail -3 gpurun_out/h2_pytest_parity.log
timeout 600 python tests/report_grad_errors.py fp16 > gpurun_out/h2_grad_errors.txt 2>&1; echo "grad report exit $?"; tail -30 gpurun_out/h2_grad_errors.txt
for round in 1 2; do for n in base default; do
  if [ $n = default ]; then unset ODVAE_LIB_PATH; else export ODVAE_LIB_PATH=$V/lib_gn_$n.so; fi
  timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['roofline']['hbm_kernels']['groupnorm']
print('[$n] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'roof', d['roofline']['frac'], 'gn', g)"
done; done | tee gpurun_out/h2_bench_ab.txt
