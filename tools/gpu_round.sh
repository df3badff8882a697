#!/bin/bash
# One GPU-box visit: parity tests, bench line, ncu launch list + tensor-pipe counters of every engine launch.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 3000 gpurun_out/bench.json
CMD="python bench.py --steps 1 --warmup 3 --no-roofline --cpu-baseline 0"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s ${NCU_SKIP:-12200} -c ${NCU_COUNT:-4100} \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 1500 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tensor.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__cycles_elapsed.max \
    --clock-control none -k regex:tc_engine -s ${ENG_SKIP:-1000} -c ${ENG_COUNT:-340} \
    --csv --log-file gpurun_out/engine_launches.csv $CMD > gpurun_out/ncu_engine.log 2>&1
echo "ncu engine exit $?"
