#!/bin/bash
# A/B of the GroupNorm backward paths on one box: two-kernel path vs the cooperative team path (hints on/off, item size).
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k groupnorm > gpurun_out/pytest_gn.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gn.log
tail -5 gpurun_out/pytest_gn.log
run() {
  name=$1; shift
  env "$@" timeout 600 python bench.py --steps 4 --warmup 3 --cpu-baseline 0 > gpurun_out/gn_ab_$name.json 2> gpurun_out/gn_ab_$name.err
  echo "$name exit $?"
  python - "$name" <<'P'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/gn_ab_{n}.json").read().strip().splitlines()[-1])
    g = d["roofline"]["hbm_kernels"]["groupnorm"]
    print(n, "img/s", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), "gn fwd", round(g["fwd"]["ms"], 2), "bwd", round(g["bwd"]["ms"], 2),
          "frac bwd", round(g["bwd"]["frac_of_hbm_peak"], 3), "sm_mhz", d["clocks"]["sm_mhz"])
except Exception as e:
    print(n, "failed", e)
P
}
run base ODVAE_GN_BWD_TEAM=0
run team ODVAE_GN_BWD_TEAM=1
run team_nohint ODVAE_GN_BWD_TEAM=1 ODVAE_GN_TEAM_HINTS=0
run team_item20 ODVAE_GN_BWD_TEAM=1 ODVAE_GN_TEAM_ITEM_MB=20
run team_item50 ODVAE_GN_BWD_TEAM=1 ODVAE_GN_TEAM_ITEM_MB=50
run team_min40 ODVAE_GN_BWD_TEAM=1 ODVAE_GN_TEAM_MIN_MB=40
