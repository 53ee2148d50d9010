#!/bin/bash
# (historical record of an experiment: the NSDG_MEVP_PF_* / NSDG_HEUN_* environment knobs it sets existed only while it ran)
# experiment: bulk L2 prefetch arm, per-warp tensor boxes (3 rows x planes) every third step
mkdir -p gpurun_out
NSDG_MEVP_PF_DIST=2 timeout 300 python tools/dyn_digest.py march_ring=3 > gpurun_out/r2y4_digest_pf.txt 2>&1; echo "pf rc=$?"; tail -1 gpurun_out/r2y4_digest_pf.txt
NSDG_MEVP_PF_DIST=3 timeout 300 python tools/dyn_digest.py march_ring=3 march_rows=7 > gpurun_out/r2y4_digest_pf7.txt 2>&1; echo "pf7 rc=$?"; tail -1 gpurun_out/r2y4_digest_pf7.txt
run() { # ring dist
  NSDG_MEVP_RING=$1 NSDG_MEVP_PF_DIST=$2 timeout 600 python bench.py --main-only --no-e2e --no-cpu > gpurun_out/r2y4_bench_r$1_d$2.json 2> gpurun_out/r2y4_bench_r$1_d$2.err; rc=$?
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2y4_bench_r$1_d$2.json"))
    print("ring=$1 dist=$2 rc=$rc", round(d["ms_per_step"],3), d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/v["launches"],4), round(v.get("frac",0),3)) for k,v in d["roofline"]["kernels"].items() if "march" in k})
except Exception as e:
    print("ring=$1 dist=$2 rc=$rc failed", e)
PY
}
run 0 0
for d in 2 3 4 6 9; do run 3 $d; done
run 0 0
NSDG_MEVP_RING=3 NSDG_MEVP_PF_DIST=3 timeout 900 ncu --set full --clock-control none -k regex:mevp_march -s 2 -c 1 -o gpurun_out/r2y4_r3 -f \
     python tools/prof_target.py 4096 8 > gpurun_out/r2y4_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r2y4_r3.ncu-rep --page raw --csv > gpurun_out/r2y4_r3_raw.csv 2>/dev/null
python - <<PY
import csv
rows=list(csv.reader(open("gpurun_out/r2y4_r3_raw.csv")))
h,u,r=rows[0],rows[1],rows[2]
for i,n in enumerate(h):
    if ("issue_stalled" in n and "per_issue_active" in n) or n in ("gpu__time_duration.sum","smsp__inst_executed.sum","smsp__issue_active.avg.pct_of_peak_sustained_active","dram__bytes_read.sum","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active","lts__t_sector_op_read_hit_rate.pct","sm__icc_request_hit_rate.pct"):
        print("  ", n, r[i], u[i])
PY
