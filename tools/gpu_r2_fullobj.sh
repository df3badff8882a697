#!/bin/bash
# ncu launch list of ONE whole batch of the yaml's own objective (LPIPS + GAN, both optimizer steps)
mkdir -p gpurun_out
export ODVAE_OVERLAP_WGRAD=0
timeout 600 python bench.py --objective full --warmup 3 --ncu-step > gpurun_out/fullobj_plain.log 2>&1 || { tail -5 gpurun_out/fullobj_plain.log; exit 1; }
timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --csv --log-file gpurun_out/fullobj_launches.csv python bench.py --objective full --warmup 3 --ncu-step > gpurun_out/fullobj_ncu.log 2>&1
echo "ncu exit $?"
python tools/ncu_pivot.py gpurun_out/fullobj_launches.csv --dram > gpurun_out/fullobj_summary.txt 2>&1
head -45 gpurun_out/fullobj_summary.txt
