#!/bin/bash
# visit: one-launch GroupNorm on the small feature maps -- tests, bench A/B (ODVAE_GN_SMALL=0 switches the forward back;
# the variant library lib_gnsmall0.so switches the backward back too)
mkdir -p gpurun_out
V=/root/repo/tools/experiments/variants
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k groupnorm > gpurun_out/sm_pytest_gn.log 2>&1; echo "pytest gn exit $?"
tail -3 gpurun_out/sm_pytest_gn.log
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_patchgan_gpu.py tests/test_graph_gpu.py tests/test_fit_gpu.py -m gpu -x -q > gpurun_out/sm_pytest_parity.log 2>&1; echo "pytest parity exit $?"
tail -3 gpurun_out/sm_pytest_parity.log
for round in 1 2; do for n in gnsmall0 default; do
  if [ $n = default ]; then unset ODVAE_LIB_PATH; export ODVAE_GN_SMALL=1; else export ODVAE_LIB_PATH=$V/lib_$n.so; export ODVAE_GN_SMALL=0; fi
  timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['roofline']['hbm_kernels']['groupnorm']
print('[$n] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'roof', round(d['roofline']['frac'],4), 'gn fwd/bwd ms', round(g['fwd']['ms'],2), round(g['bwd']['ms'],2), 'host', round(d.get('host_enqueue_ms_per_step'),1), 'launches', d.get('gpu_launches'))"
done; done | tee gpurun_out/sm_bench_ab.txt
