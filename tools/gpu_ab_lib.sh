#!/bin/bash
# A/B of two builds of libodvae_b200.so on ONE box (boxes differ by +-3 % under the power cap):
# generative-detection_b200/odvae_b200/libodvae_b200_old.so against the in-tree build, alternating.
mkdir -p gpurun_out
OLD=$PWD/generative-detection_b200/odvae_b200/libodvae_b200_old.so
for i in 1 2; do
  for v in old new; do
    if [ $v = old ]; then export ODVAE_LIB_PATH=$OLD; else unset ODVAE_LIB_PATH; fi
    timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/ab_$v$i.json 2> gpurun_out/ab_$v$i.err
    python -c "
import json;d=json.loads(open('gpurun_out/ab_$v$i.json').read().strip().splitlines()[-1]);print('$v$i',round(d['value'],1),round(d['ms_per_step'],2),round(d['roofline']['frac'],4),d['clocks']['sm_mhz'], {k:round(v['ms'],2) for k,v in d['roofline']['by_call'].items()})"
  done
done
