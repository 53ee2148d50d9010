#!/bin/bash
# (historical record of an experiment: the NSDG_MEVP_PF_* / NSDG_HEUN_* environment knobs it sets existed only while it ran)
# why is the bulk L2 prefetch arm slower?  DRAM bytes / L2 hit rate of one launch, distance 2 and 8
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sector_op_read_hit_rate.pct,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_tex_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_membar_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
for cfg in "0 0" "3 2" "3 8"; do
  set -- $cfg
  NSDG_MEVP_RING=$1 NSDG_MEVP_PF_DIST=$2 timeout 600 ncu --metrics $M --clock-control none -k regex:mevp_march -s 2 -c 1 --csv --log-file gpurun_out/r2z_ncu_r$1_d$2.csv \
     python tools/prof_target.py 4096 8 > gpurun_out/r2z_ncu_r$1_d$2.log 2>&1; echo "ncu ring=$1 dist=$2 rc=$?"
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r2z_ncu_r$1_d$2.csv")) if len(r)>10]
h=rows[0]; mi=h.index("Metric Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
for r in rows[1:]: print("  ring=$1 dist=$2", r[mi], r[vi], r[ui])
PY
done
