#!/bin/bash
# Launch list (duration + DRAM bytes per launch) of ONE train step, bracketed by cudaProfilerStart/Stop in bench.py --ncu-step.
mkdir -p gpurun_out
CMD="python bench.py --warmup 3 --ncu-step"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
