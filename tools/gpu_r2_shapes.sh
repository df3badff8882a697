#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 --extras 0 --cpu-baseline 0 --shapes gpurun_out/shapes.json > gpurun_out/shapes_bench.json 2> gpurun_out/shapes.err; echo "exit $?"
tail -c 600 gpurun_out/shapes_bench.json
