#!/bin/bash
# same-box A/B of library variants built into tools/experiments/variants/lib_<name>.so (ODVAE_LIB_PATH picks one):
#   bash tools/gpu_ab.sh name1 name2 ...   -> conv probe, attention probe and a 10-step bench per variant, two rounds
V=/root/repo/tools/experiments/variants
for round in 1 2; do for n in "$@"; do
export ODVAE_LIB_PATH=$V/lib_$n.so
timeout 300 python tools/conv_probe.py 3 2>&1 | tail -1 | sed "s/^/[$n] /"
timeout 300 python tools/attn_probe.py 64 fp16 | sed "s/^/[$n] /"
timeout 600 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('[$n] bench', round(d['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'])"
done; done
