#!/bin/bash
# L2 prefetch arm again, now that the loop body (31.1 KB with it) fits the 32 KB instruction-cache level
mkdir -p gpurun_out
for r in 0 3 0 3; do
NSDG_MEVP_RING=$r timeout 600 python bench.py --main-only --no-e2e --no-cpu > gpurun_out/r2ah_bench_r$r.json 2> gpurun_out/r2ah_bench_r$r.err; echo "bench ring=$r rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2ah_bench_r$r.json"))
print("ring=$r", round(d["ms_per_step"],3), d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/v["launches"],4), round(v.get("frac",0),3)) for k,v in d["roofline"]["kernels"].items() if "mevp" in k})
PY
done
M=gpu__time_duration.sum,dram__bytes_read.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__icc_request_hit_rate.pct,lts__t_sector_op_read_hit_rate.pct,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
NSDG_MEVP_RING=3 timeout 600 ncu --metrics $M --clock-control none -k regex:mevp_march -s 2 -c 1 --csv --log-file gpurun_out/r2ah_ncu.csv python tools/prof_target.py 4096 8 > gpurun_out/r2ah_ncu.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r2ah_ncu.csv")) if len(r)>10]
h=rows[0]; mi=h.index("Metric Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
for r in rows[1:]: print("  ", r[mi], r[vi], r[ui])
PY
