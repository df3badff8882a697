#!/bin/bash
for bs in 0 1; do for dbg in 0 1 2; do
ODVAE_DBG=$dbg ODVAE_GEMM_BSTAT=$bs timeout 300 python tools/attn_probe.py 64 fp16 | sed "s/^/dbg=$dbg /"
done; done
