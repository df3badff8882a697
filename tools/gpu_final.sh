#!/bin/bash
# Round-end evidence for HEAD: bench line (+ per-shape table) and the launch list of one step with DRAM bytes.
set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 --shapes gpurun_out/shapes.json > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 2500 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
CMD="python bench.py --warmup 3 --ncu-step"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
