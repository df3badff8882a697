#!/bin/bash
# visit: GroupNorm-backward sums fused into the dgrad epilogue -- tests, parity, bench A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_engine_gpu.py -m gpu -x -q -k "fused_groupnorm_backward or halo" > gpurun_out/gnb_pytest.log 2>&1; echo "pytest engine exit $?"
tail -4 gpurun_out/gnb_pytest.log
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_graph_gpu.py tests/test_fit_gpu.py -m gpu -x -q > gpurun_out/gnb_pytest_parity.log 2>&1; echo "pytest parity exit $?"
tail -3 gpurun_out/gnb_pytest_parity.log
for round in 1 2; do for m in 0 1; do
  ODVAE_GN_BWD_FUSE=$m timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['roofline']['hbm_kernels']['groupnorm']
print('[fuse=$m] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'roof', round(d['roofline']['frac'],4), 'gn fwd/bwd ms', round(g['fwd']['ms'],2), round(g['bwd']['ms'],2), 'bwd frac', round(g['bwd']['frac_of_hbm_peak'],3), 'launches', d['gpu_launches']//10)"
done; done | tee gpurun_out/gnb_bench_ab.txt
