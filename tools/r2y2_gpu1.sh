#!/bin/bash
# (historical record of an experiment: the NSDG_MEVP_PF_* / NSDG_HEUN_* environment knobs it sets existed only while it ran)
# experiment: bulk L2 prefetch arm, elected-lane CTA-wide variant (pf_cta = 2)
mkdir -p gpurun_out
NSDG_MEVP_PF_DIST=2 NSDG_MEVP_PF_CTA=2 timeout 300 python tools/dyn_digest.py march_ring=3 > gpurun_out/r2y2_digest_pf.txt 2>&1; echo "pf rc=$?"; tail -1 gpurun_out/r2y2_digest_pf.txt
NSDG_MEVP_PF_DIST=3 NSDG_MEVP_PF_CTA=2 timeout 300 python tools/dyn_digest.py march_ring=3 march_rows=7 > gpurun_out/r2y2_digest_pf7.txt 2>&1; echo "pf7 rc=$?"; tail -1 gpurun_out/r2y2_digest_pf7.txt
run() { # ring dist cta
  NSDG_MEVP_RING=$1 NSDG_MEVP_PF_DIST=$2 NSDG_MEVP_PF_CTA=$3 timeout 600 python bench.py --main-only --no-e2e --no-cpu > gpurun_out/r2y2_bench_r$1_d$2_c$3.json 2> gpurun_out/r2y2_bench_r$1_d$2_c$3.err; rc=$?
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2y2_bench_r$1_d$2_c$3.json"))
    print("ring=$1 dist=$2 cta=$3 rc=$rc", round(d["ms_per_step"],3), d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/v["launches"],4), round(v.get("frac",0),3)) for k,v in d["roofline"]["kernels"].items() if "march" in k})
except Exception as e:
    print("ring=$1 dist=$2 cta=$3 rc=$rc failed", e)
PY
}
run 0 0 0
for d in 1 2 3 4; do run 3 $d 2; done
run 0 0 0
