#!/bin/bash
mkdir -p gpurun_out
for v in base v4 v5 v6 v7 v8; do
  if [ $v = base ]; then unset ODVAE_LIB_PATH; else export ODVAE_LIB_PATH=$PWD/generative-detection_b200/odvae_b200/libodvae_gn_$v.so; fi
  echo "== $v"
  timeout 300 python tools/gn_bench.py 5 2>&1 | tail -4
done
