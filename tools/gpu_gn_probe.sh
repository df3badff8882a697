#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k groupnorm > gpurun_out/pytest_gn.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gn.log
tail -5 gpurun_out/pytest_gn.log
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct
for v in "base ODVAE_GN_BWD_TEAM=0" "team2 ODVAE_GN_TEAMS=2" "team2nohint ODVAE_GN_TEAMS=2 ODVAE_GN_TEAM_HINTS=0" "team1 ODVAE_GN_TEAMS=1" "team2_add ODVAE_GN_TEAMS=2 PADD=1" "base_add ODVAE_GN_BWD_TEAM=0 PADD=1" "team2_128 ODVAE_GN_TEAMS=2 PH=128 PB=256" "team2_64 ODVAE_GN_TEAMS=2 PH=64 PB=256 PC=256"  "base_64 ODVAE_GN_BWD_TEAM=0 PH=64 PB=256 PC=256"; do
  set -- $v; name=$1; shift
  env "$@" timeout 300 python tools/gn_probe.py 4 > gpurun_out/gnp_$name.log 2>&1; echo "$name exit $?"; tail -2 gpurun_out/gnp_$name.log
done
for v in "team2 ODVAE_GN_TEAMS=2" "team2_add ODVAE_GN_TEAMS=2 PADD=1"; do
  set -- $v; name=$1; shift
  env "$@" timeout 600 ncu --metrics $M --clock-control none -k regex:gn_ --csv --log-file gpurun_out/gnp_$name.csv python tools/gn_probe.py 2 > gpurun_out/gnp_${name}_ncu.log 2>&1
  echo "ncu $name exit $?"
done
