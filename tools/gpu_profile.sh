#!/bin/bash
# Launch list of ONE train step (cudaProfilerStart/Stop bracket inside bench.py --ncu-step) + per-shape engine table.
set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 --cpu-baseline 0 --shapes gpurun_out/shapes.json > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 1500 gpurun_out/bench.json
CMD="python bench.py --warmup 3 --ncu-step"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
