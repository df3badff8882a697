#!/bin/bash
# One GPU-box visit: parity tests, full bench line (+ per-shape table), launch list of one step with DRAM bytes,
# and one `ncu --set full` capture of the dominant conv shape (fwd / fwd+res / dgrad / wgrad of 128->128 3x3 @256^2).
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 --shapes gpurun_out/shapes.json > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 2500 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
CMD="python bench.py --warmup 3 --ncu-step"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 300 python tools/conv_probe.py 3 > gpurun_out/conv_probe.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_engine -c 4 -o gpurun_out/conv_full -f \
    python tools/conv_probe.py 1 > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"; cat gpurun_out/conv_probe.log
