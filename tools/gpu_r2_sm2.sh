#!/bin/bash
mkdir -p gpurun_out
V=/root/repo/tools/experiments/variants
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k groupnorm > gpurun_out/sm2_pytest_gn.log 2>&1; echo "pytest gn exit $?"
tail -3 gpurun_out/sm2_pytest_gn.log
for n in gnprev default gnprev default; do
  if [ $n = default ]; then unset ODVAE_LIB_PATH; else export ODVAE_LIB_PATH=$V/lib_$n.so; fi
  echo "== gn_bench $n"; timeout 300 python tools/gn_bench.py 5 2>&1 | tail -1
done | tee gpurun_out/sm2_gn_bench.txt
unset ODVAE_LIB_PATH
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_graph_gpu.py -m gpu -x -q > gpurun_out/sm2_pytest_parity.log 2>&1; echo "pytest parity exit $?"
tail -2 gpurun_out/sm2_pytest_parity.log
for round in 1 2; do for n in gnprev default; do
  if [ $n = default ]; then unset ODVAE_LIB_PATH; else export ODVAE_LIB_PATH=$V/lib_$n.so; fi
  timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['roofline']['hbm_kernels']['groupnorm']
print('[$n] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'roof', round(d['roofline']['frac'],4), 'gn fwd/bwd ms', round(g['fwd']['ms'],2), round(g['bwd']['ms'],2))"
done; done | tee gpurun_out/sm2_bench_ab.txt
