#!/bin/bash
# visit: specialised swapped-conv epilogue + satfinite packs + GroupNorm dgamma/dbeta sums inside the apply kernels
mkdir -p gpurun_out
V=/root/repo/tools/experiments/variants
timeout 1200 python -m pytest tests/test_engine_gpu.py tests/test_kernels_gpu.py -m gpu -x -q > gpurun_out/fe_pytest_k.log 2>&1; echo "pytest kernels exit $?"
tail -3 gpurun_out/fe_pytest_k.log
for n in default fe0 default fe0; do
  if [ $n = default ]; then unset ODVAE_LIB_PATH; else export ODVAE_LIB_PATH=$V/lib_$n.so; fi
  echo "== conv_probe $n"; timeout 300 python tools/conv_probe.py 3 2>&1 | tail -2
done | tee gpurun_out/fe_conv_probe.txt
unset ODVAE_LIB_PATH
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_patchgan_gpu.py tests/test_graph_gpu.py tests/test_fit_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/fe_pytest_parity.log 2>&1; echo "pytest parity exit $?"
tail -3 gpurun_out/fe_pytest_parity.log
for round in 1 2; do for n in gn_base fe0 default; do
  if [ $n = default ]; then unset ODVAE_LIB_PATH; else export ODVAE_LIB_PATH=$V/lib_$n.so; fi
  timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['roofline']['hbm_kernels']['groupnorm']
print('[$n] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'roof', round(d['roofline']['frac'],4), 'gn fwd/bwd ms', round(g['fwd']['ms'],2), round(g['bwd']['ms'],2), 'host', d.get('host_enqueue_ms_per_step'), 'launches', d.get('gpu_launches'))"
done; done | tee gpurun_out/fe_bench_ab.txt
