#!/bin/bash
# visit: W = 128 halo tiles (halo == 2) -- tests, conv probe at 128x128 A/B, parity, bench A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_engine_gpu.py -m gpu -x -q -k "halo or conv_fwd_dgrad or gnstats or fused_groupnorm" > gpurun_out/h128_pytest.log 2>&1; echo "pytest engine exit $?"
tail -4 gpurun_out/h128_pytest.log
for m in 1 0 1 0; do echo "== conv_probe 128x128 HALO2=$m"; PH=128 ODVAE_CONV_HALO2=$m timeout 300 python tools/conv_probe.py 3 2>&1 | tail -2; done | tee gpurun_out/h128_conv_probe.txt
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_graph_gpu.py tests/test_kernels_gpu.py -m gpu -x -q > gpurun_out/h128_pytest_parity.log 2>&1; echo "pytest parity exit $?"
tail -3 gpurun_out/h128_pytest_parity.log
for round in 1 2; do for m in 0 1; do
  ODVAE_CONV_HALO2=$m timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('[halo2=$m] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'roof', round(d['roofline']['frac'],4))"
done; done | tee gpurun_out/h128_bench_ab.txt
