#!/bin/bash
# (historical record of an experiment: the NSDG_MEVP_PF_* / NSDG_HEUN_* environment knobs it sets existed only while it ran)
# all stall reasons of the marching kernel with and without the elected-lane bulk L2 prefetch
mkdir -p gpurun_out
for cfg in "0 0 0" "3 2 2"; do
  set -- $cfg
  NSDG_MEVP_RING=$1 NSDG_MEVP_PF_DIST=$2 NSDG_MEVP_PF_CTA=$3 timeout 900 ncu --set full --clock-control none -k regex:mevp_march -s 2 -c 1 -o gpurun_out/r2z2_r$1_c$3 -f \
     python tools/prof_target.py 4096 8 > gpurun_out/r2z2_ncu_r$1_c$3.log 2>&1; echo "ncu ring=$1 rc=$?"
  ncu -i gpurun_out/r2z2_r$1_c$3.ncu-rep --page raw --csv > gpurun_out/r2z2_r$1_c$3_raw.csv 2>/dev/null
  python - <<PY
import csv
rows=list(csv.reader(open("gpurun_out/r2z2_r$1_c$3_raw.csv")))
h,u,r=rows[0],rows[1],rows[2]
for i,n in enumerate(h):
    if ("issue_stalled" in n and "per_issue_active" in n) or n in ("gpu__time_duration.sum","smsp__inst_executed.sum","smsp__issue_active.avg.pct_of_peak_sustained_active","dram__bytes_read.sum","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active") or "icc" in n or "instruction_cache" in n:
        print("  ring=$1", n, r[i], u[i])
PY
done
