#!/bin/bash
mkdir -p gpurun_out
V=/root/repo/tools/experiments/variants
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -x -q -k "fused_groupnorm_backward" 2>&1 | tail -2
for round in 1 2 3; do for n in nopf default; do
  if [ $n = default ]; then unset ODVAE_LIB_PATH; else export ODVAE_LIB_PATH=$V/lib_$n.so; fi
  timeout 600 python bench.py --steps 10 --warmup 3 --extras 0 --cpu-baseline 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('[$n] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], 'dgrad+gnbwd ms', round(d['roofline']['by_call']['conv2d_dgrad_gnbwd']['ms'],2))"
done; done | tee gpurun_out/pf_bench_ab.txt
