#!/bin/bash
# fused two-stage dG1 transport step (transport_heun_march_kernel): parity tests, cone bench with and without
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dynamics.py -m gpu -x -q -k "transport or variants or cone or mass or graph" > gpurun_out/r2aa_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2aa_pytest.log
for fz in 1 0; do
  NSDG_TRANSPORT_FUSED=$fz timeout 600 python bench.py --workload cone --no-e2e > gpurun_out/r2aa_cone_fused$fz.json 2> gpurun_out/r2aa_cone_fused$fz.err; echo "cone fused=$fz rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2aa_cone_fused$fz.json"))
r=d["roofline"]
print("fused=$fz", d["value"], "us/step", round(d["ms_per_step"]*1e3,2), "launches", d["gpu_launches"], r["kernel"], "avg us", round(r["avg_launch_ms"]*1e3,2), "clk", d["clocks"]["sm_mhz"], "drift", d["mass_drift_rel"])
PY
done
