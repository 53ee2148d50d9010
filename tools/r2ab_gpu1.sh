#!/bin/bash
# (historical record of an experiment: the NSDG_MEVP_PF_* / NSDG_HEUN_* environment knobs it sets existed only while it ran)
# sweep: resident CTAs per SM / band height of the fused Heun kernel on the 1024^2 cone
mkdir -p gpurun_out
run() {
  NSDG_HEUN_MINB=$1 NSDG_HEUN_ROWS=$2 timeout 600 python bench.py --workload cone --no-e2e --no-cpu > gpurun_out/r2ab_cone_m$1_r$2.json 2> gpurun_out/r2ab_cone_m$1_r$2.err; rc=$?
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2ab_cone_m$1_r$2.json"))
    print("minb=$1 rows=$2", round(d["value"]/1e10,3), "us/step", round(d["ms_per_step"]*1e3,2), "host us", round(d["host_enqueue_us_per_step"],2), "clk", d["clocks"]["sm_mhz"])
except Exception as e: print("minb=$1 rows=$2 rc=$rc failed", e)
PY
}
for m in 3 4 5 6; do run $m 0; done
for r in 12 14 16 21 24 28 35; do run 4 $r; done
for r in 14 21; do run 5 $r; done
for r in 28; do run 3 $r; done
