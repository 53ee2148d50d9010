#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/dyn_digest.py > gpurun_out/r2w_digest_default.txt 2>&1; echo "default rc=$?"; tail -1 gpurun_out/r2w_digest_default.txt
timeout 300 python tools/dyn_digest.py march_ring=2 > gpurun_out/r2w_digest_lring.txt 2>&1; echo "lring rc=$?"; tail -2 gpurun_out/r2w_digest_lring.txt
timeout 300 python tools/dyn_digest.py march_ring=2 march_rows=5 > gpurun_out/r2w_digest_lring5.txt 2>&1; echo "lring5 rc=$?"; tail -1 gpurun_out/r2w_digest_lring5.txt
for r in 0 2; do
NSDG_MEVP_RING=$r timeout 600 python bench.py --main-only --no-e2e --no-cpu > gpurun_out/r2w_bench_ring$r.json 2> gpurun_out/r2w_bench_ring$r.err; echo "bench ring=$r rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2w_bench_ring$r.json"))
print("ring=$r", d["ms_per_step"], d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/v["launches"],4), round(v.get("frac",0),3)) for k,v in d["roofline"]["kernels"].items()})
PY
done
