#!/bin/bash
# marching sub-cycle kernel with the plane row carried from step to step (no modulo in the loop body: 29.7 KB instead of 33.1 KB)
mkdir -p gpurun_out
timeout 300 python tools/dyn_digest.py > gpurun_out/r2ag_digest.txt 2>&1; echo "digest rc=$?"; tail -1 gpurun_out/r2ag_digest.txt
timeout 300 python tools/dyn_digest.py march_rows=5 > gpurun_out/r2ag_digest5.txt 2>&1; echo "digest5 rc=$?"; tail -1 gpurun_out/r2ag_digest5.txt
timeout 300 python tools/dyn_digest.py march_ring=3 march_rows=7 > gpurun_out/r2ag_digest_pf.txt 2>&1; echo "digestpf rc=$?"; tail -1 gpurun_out/r2ag_digest_pf.txt
for i in 1 2; do
timeout 600 python bench.py --main-only --no-e2e --no-cpu > gpurun_out/r2ag_bench$i.json 2> gpurun_out/r2ag_bench$i.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2ag_bench$i.json"))
print("run $i", round(d["ms_per_step"],3), d["value"], d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/v["launches"],4), v["launches"], round(v.get("frac",0),3)) for k,v in d["roofline"]["kernels"].items()})
PY
done
M=gpu__time_duration.sum,dram__bytes_read.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__icc_request_hit_rate.pct,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active
timeout 600 ncu --metrics $M --clock-control none -k regex:mevp_march -s 2 -c 1 --csv --log-file gpurun_out/r2ag_ncu.csv python tools/prof_target.py 4096 8 > gpurun_out/r2ag_ncu.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r2ag_ncu.csv")) if len(r)>10]
h=rows[0]; mi=h.index("Metric Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
for r in rows[1:]: print("  ", r[mi], r[vi], r[ui])
PY
