#!/bin/bash
# BASELINE.json configs[2] (encode-only sweep, batch 1..512) and configs[4] (512x512 crops), one JSON line each.
mkdir -p gpurun_out
: > gpurun_out/encode_sweep.jsonl
for b in 1 2 4 8 16 32 64 128 256 512; do
  timeout 300 python bench.py --workload encode --batch $b --steps 10 --warmup 3 >> gpurun_out/encode_sweep.jsonl 2>> gpurun_out/encode_sweep.err || echo "encode batch $b failed"
done
cut -c1-330 gpurun_out/encode_sweep.jsonl
timeout 600 python bench.py --workload recon --res 512 --batch 8 --steps 5 --warmup 3 --cpu-baseline 0 > gpurun_out/bench_res512.json 2> gpurun_out/bench_res512.err; echo "res512 exit $?"
cut -c1-600 gpurun_out/bench_res512.json; tail -3 gpurun_out/bench_res512.err
