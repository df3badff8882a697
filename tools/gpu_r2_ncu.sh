#!/bin/bash
# round 2 ncu evidence: (1) launch list of one train step of the round-end code with DRAM bytes, (2) --set full captures of the
# GroupNorm kernels and of the engine instances (conv swapped-role fwd / wgrad / phase up-conv) inside the same step
set -x
mkdir -p gpurun_out
export ODVAE_OVERLAP_WGRAD=0   # ncu serialises kernels anyway; keep every launch on one stream
CMD="python bench.py --warmup 3 --ncu-step"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 || exit 1
timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
python tools/ncu_pivot.py gpurun_out/launches.csv --dram > gpurun_out/launches_summary.txt 2>&1
timeout 1200 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"gn_apply|gn_bwd_apply|gn_reduce" -s 8 -c 6 \
    -o gpurun_out/gn_full -f $CMD > gpurun_out/ncu_gn.log 2>&1
echo "ncu gn exit $?"
timeout 1500 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"tc_engine" -s 30 -c 10 \
    -o gpurun_out/engine_full -f $CMD > gpurun_out/ncu_engine.log 2>&1
echo "ncu engine exit $?"
ls -la gpurun_out/*.ncu-rep
head -30 gpurun_out/launches_summary.txt
