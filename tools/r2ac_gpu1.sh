#!/bin/bash
# stages 1 + 2 of the dG2 SSP-RK3 step in one launch: parity tests, north-star step with and without
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_dynamics.py -m gpu -x -q > gpurun_out/r2ac_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2ac_pytest.log
for fz in 1 0; do
  NSDG_TRANSPORT_FUSED=$fz timeout 600 python bench.py --main-only --no-e2e --no-cpu > gpurun_out/r2ac_bench_fused$fz.json 2> gpurun_out/r2ac_bench_fused$fz.err; echo "bench fused=$fz rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2ac_bench_fused$fz.json"))
print("fused=$fz", round(d["ms_per_step"],3), d["value"], d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/v["launches"],4), v["launches"], round(v.get("frac",0),3)) for k,v in d["roofline"]["kernels"].items()})
print("   whole step", d["roofline"]["whole_step"])
PY
done
NSDG_TRANSPORT_FUSED=1 timeout 600 python bench.py --workload cone --no-e2e --no-cpu > gpurun_out/r2ac_cone.json 2> gpurun_out/r2ac_cone.err; echo "cone rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2ac_cone.json')); print('cone', d['value'], round(d['ms_per_step']*1e3,2), d['launch_mode'])"
